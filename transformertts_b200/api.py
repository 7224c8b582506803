"""Drop-in for the reference's `model.TransformerTTS` (model/model.py:138-394) on B200.

Same constructor kwargs, same `forward` / `inference` signatures and output dicts, same `state_dict`
key names, so `lightning_module.py:30,42,154` can use it unchanged for the no-grad / validation /
inference passes.  All arithmetic runs in the hand-written CUDA library (libtts_b200.so) through the C ABI
of include/tts_b200.h; PyTorch only owns the tensors.  There is no CPU path and no autograd: calling
`forward` in training mode raises.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional

import torch
import torch.nn as nn

from . import _ext
from .weights import D, FFN, KS, MAX_PE, N_MELS, PackedWeights

FLAG_WANT_ALIGN, FLAG_ENC_MASK, FLAG_SERIAL, FLAG_MEL_BF16 = 1, 2, 4, 8


# Static outputs of graph replays that another stream is still reading (data.OutputFetcher's device -> host copies):
# (device index, data_ptr of pred_melspec) -> events the next replay into that buffer must wait for.  record_stream()
# cannot protect these tensors -- they live in a graph's private pool and the replay overwrites them regardless.
_STATIC_OUTPUT_READERS: Dict[tuple, list] = {}


def note_output_read(pred: torch.Tensor, event: "torch.cuda.Event") -> bool:
    """Tell the module that `event` (recorded on the reader's stream) marks the end of a cross-stream read of the forward
    outputs whose `pred_melspec` is `pred`.  No-op (False) unless that tensor is a static graph output."""
    readers = _STATIC_OUTPUT_READERS.get((pred.device.index, pred.data_ptr()))
    if readers is None:
        return False
    readers.append(event)
    return True


class _GraphEntry:
    """One captured forward per output set, sharing static INPUT buffers the caller's tensors are copied into."""
    __slots__ = ("inputs", "sets", "turn", "weights", "launches")


class _Holder(nn.Module):
    """Plain container node so that parameter names reproduce the reference's dotted keys."""


def _pe_table() -> torch.Tensor:
    # model/model.py:80-86
    pos = torch.arange(0, MAX_PE, dtype=torch.float32).unsqueeze(1)
    div = torch.exp(torch.arange(0, D, 2, dtype=torch.float32) * (-math.log(10000.0)) / D)
    pe = torch.zeros(MAX_PE, D)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe


def _reference_layout_init(n_phon: int) -> Dict[str, torch.Tensor]:
    """Fresh tensors under the reference's key names with the reference's init laws
    (model/module.py:21-25,47-50 Xavier-normal with activation gain, zero bias; torch defaults elsewhere)."""
    sd: Dict[str, torch.Tensor] = {}
    relu_gain, tanh_gain = math.sqrt(2.0), 5.0 / 3.0

    def xavier(shape, gain):
        t = torch.empty(shape)
        nn.init.xavier_normal_(t, gain=gain)
        return t

    def conv_bn(prefix, cin, cout, gain):
        sd[f"{prefix}.conv.weight"] = xavier((cout, cin, KS), gain)
        sd[f"{prefix}.conv.bias"] = torch.zeros(cout)
        sd[f"{prefix}.bn.weight"] = torch.ones(cout)
        sd[f"{prefix}.bn.bias"] = torch.zeros(cout)
        sd[f"{prefix}.bn.running_mean"] = torch.zeros(cout)
        sd[f"{prefix}.bn.running_var"] = torch.ones(cout)
        sd[f"{prefix}.bn.num_batches_tracked"] = torch.zeros((), dtype=torch.int64)

    def linear_norm(prefix, cin, cout):
        sd[f"{prefix}.linear.weight"] = xavier((cout, cin), relu_gain)
        sd[f"{prefix}.linear.bias"] = torch.zeros(cout)

    def mha(prefix):
        w = torch.empty(3 * D, D)
        nn.init.xavier_uniform_(w)
        sd[f"{prefix}.in_proj_weight"] = w
        sd[f"{prefix}.in_proj_bias"] = torch.zeros(3 * D)
        lin = nn.Linear(D, D)
        sd[f"{prefix}.out_proj.weight"] = lin.weight.detach().clone()
        sd[f"{prefix}.out_proj.bias"] = torch.zeros(D)

    def ffn_ln(prefix, n_norm):
        l1, l2 = nn.Linear(D, FFN), nn.Linear(FFN, D)
        sd[f"{prefix}.linear1.weight"], sd[f"{prefix}.linear1.bias"] = l1.weight.detach().clone(), l1.bias.detach().clone()
        sd[f"{prefix}.linear2.weight"], sd[f"{prefix}.linear2.bias"] = l2.weight.detach().clone(), l2.bias.detach().clone()
        for i in range(1, n_norm + 1):
            sd[f"{prefix}.norm{i}.weight"] = torch.ones(D)
            sd[f"{prefix}.norm{i}.bias"] = torch.zeros(D)

    sd["emb.weight"] = torch.randn(n_phon, D)
    for i in range(3):
        conv_bn(f"enc_prenet.layers.{2 * i}", D, D, relu_gain)
    linear_norm("enc_prenet.linear", D, D)
    linear_norm("dec_prenet.linear1", N_MELS, D)
    linear_norm("dec_prenet.linear2", D, D)
    sd["pe.alpha"] = torch.ones(1)
    sd["pe.pe"] = _pe_table()
    for l in range(3):
        mha(f"encoder.layers.{l}.self_attn")
        ffn_ln(f"encoder.layers.{l}", 2)
    for l in range(3):
        mha(f"decoder.layers.{l}.self_attn")
        mha(f"decoder.layers.{l}.multihead_attn")
        ffn_ln(f"decoder.layers.{l}", 3)
    chans = [N_MELS, D, D, D, D, N_MELS]
    for i in range(5):
        conv_bn(f"postnet.layers.{3 * i}", chans[i], chans[i + 1], tanh_gain)
    linear_norm("linear1", D, N_MELS)
    linear_norm("linear2", D, 1)
    return sd


_BUFFER_SUFFIXES = ("running_mean", "running_var", "num_batches_tracked", "pe.pe")


class TransformerTTS(nn.Module):
    """B200 replacement for model/model.py:138 `TransformerTTS` (eval-mode forward and AR inference)."""

    def __init__(self, encoder_prenet_n_layers=3, encoder_prenet_in_channel=256, encoder_prenet_out_channel=256,
                 encoder_prenet_kernel_size=5, encoder_prenet_dropout=0.5, encoder_n_layers=3, encoder_n_head=4,
                 encoder_d_ffn=1024, encoder_dropout=0.1, decoder_n_layers=3, decoder_n_head=4, decoder_d_ffn=1024,
                 decoder_dropout=0.1, postnet_n_layers=5, postnet_kernel_size=5, postnet_dropout=0.5, d_model=256,
                 n_phon=100, n_mels=80, device="cuda"):
        super().__init__()
        got = dict(encoder_prenet_n_layers=encoder_prenet_n_layers, encoder_prenet_in_channel=encoder_prenet_in_channel,
                   encoder_prenet_out_channel=encoder_prenet_out_channel,
                   encoder_prenet_kernel_size=encoder_prenet_kernel_size, encoder_n_layers=encoder_n_layers,
                   encoder_n_head=encoder_n_head, encoder_d_ffn=encoder_d_ffn, decoder_n_layers=decoder_n_layers,
                   decoder_n_head=decoder_n_head, decoder_d_ffn=decoder_d_ffn, postnet_n_layers=postnet_n_layers,
                   postnet_kernel_size=postnet_kernel_size, d_model=d_model, n_mels=n_mels)
        want = dict(encoder_prenet_n_layers=3, encoder_prenet_in_channel=256, encoder_prenet_out_channel=256,
                    encoder_prenet_kernel_size=5, encoder_n_layers=3, encoder_n_head=4, encoder_d_ffn=1024,
                    decoder_n_layers=3, decoder_n_head=4, decoder_d_ffn=1024, postnet_n_layers=5,
                    postnet_kernel_size=5, d_model=256, n_mels=80)
        bad = {k: v for k, v in got.items() if want[k] != v}
        if bad:
            raise ValueError(f"the sm_100a kernels are specialised to config.yaml:25-42; unsupported values: {bad}")
        self.device = torch.device(device)
        self.n_phon = n_phon
        self.serial_encoder = False   # True: keep the encoder branch on the caller's stream (no side stream)
        self.use_cuda_graph = False   # True: replay the forward as a CUDA graph (static output buffers; _forward_graph)
        self.graph_output_sets = 2    # static output sets a graph entry rotates through (next call never overwrites the last result)
        self._graphs = {}
        self.graph_launches = 0       # kernel launches executed by graph replays (tts_launch_count sees only captures)
        self.strict_lens = False      # True: reproduce the reference's Tp == max(lens) assert (costs a host sync)
        for key, t in _reference_layout_init(n_phon).items():
            self._register(key, t.to(self.device))
        self._packed: Optional[PackedWeights] = None
        self._packed_fp = None
        self._fp_tensors = None
        self._decode_capacity = None
        self._warned_capacity = False
        self._ws = None
        self._scratch_bufs = {}
        self.register_load_state_dict_post_hook(lambda module, incompatible: module.refresh_weights())

    # ------------------------------------------------------------------ parameter tree
    def _register(self, key: str, t: torch.Tensor):
        node = self
        parts = key.split(".")
        for p in parts[:-1]:
            if not hasattr(node, p):
                node.add_module(p, _Holder())
            node = getattr(node, p)
        if key.endswith(_BUFFER_SUFFIXES):
            node.register_buffer(parts[-1], t)
        else:
            node.register_parameter(parts[-1], nn.Parameter(t, requires_grad=t.is_floating_point()))

    @classmethod
    def from_reference(cls, module_or_state_dict, device="cuda") -> "TransformerTTS":
        """Build from a reference `model.TransformerTTS` instance or its state_dict (same key names)."""
        sd = module_or_state_dict if isinstance(module_or_state_dict, dict) else module_or_state_dict.state_dict()
        m = cls(n_phon=sd["emb.weight"].shape[0], device=device)
        m.load_state_dict(sd, strict=True)
        return m.eval()

    def refresh_weights(self):
        """Drop the packed (bf16 / BN-folded) device copies; they are rebuilt from the parameters on next use."""
        self._packed, self._fp_tensors = None, None

    def _fingerprint(self):
        """(data_ptr, in-place version counter) of every parameter and buffer: changes whenever a tensor is replaced,
        moved or updated in place (optimizer step, `p.fill_()` under no_grad, BN statistic edits).  Writes through
        `p.data` bypass torch's version counter: call `refresh_weights()` after those."""
        if self._fp_tensors is None:      # the module-tree walk is the slow part: keep the flat list
            self._fp_tensors = list(self.parameters()) + list(self.buffers())
        return tuple((t.data_ptr(), t._version) for t in self._fp_tensors)

    def _weights(self) -> PackedWeights:
        fp = self._fingerprint()
        if self._packed is None or self._packed_fp != fp:
            self._packed = PackedWeights(self.state_dict(), self.device)
            self._packed_fp = fp
        return self._packed

    def _apply(self, fn, *args, **kwargs):
        """`.to()` / `.cuda()` / `.float()`: keep `self.device` in step with the parameters and drop the packed copies."""
        out = super()._apply(fn, *args, **kwargs)
        dev = self.emb.weight.device
        if dev.type == "cuda":
            self.device = dev
        self._packed, self._ws, self._fp_tensors = None, None, None
        self._scratch_bufs = {}
        self._drop_graphs()
        return out

    def _drop_graphs(self):
        for e in self._graphs.values():
            for _, o in e.sets:
                _STATIC_OUTPUT_READERS.pop((o["pred_melspec"].device.index, o["pred_melspec"].data_ptr()), None)
        self._graphs = {}

    def _workspace(self, B, Tp, Tm):
        need = _ext.load().tts_workspace_bytes(B, Tp, Tm) + 1024
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        base = self._ws.data_ptr()
        aligned = (base + 1023) & ~1023
        return C.c_void_p(aligned), self._ws.numel() - (aligned - base)

    def _scratch(self, name, numel, dtype):
        """Grow-only per-module scratch for inference(); like the workspace it is shared by all calls on this module, which
        are ordered on one stream.  The returned results are always fresh tensors (copies of the produced prefix)."""
        t = self._scratch_bufs.get(name)
        if t is None or t.numel() < numel or t.device != self.device:
            t = torch.empty(numel, dtype=dtype, device=self.device)
            self._scratch_bufs[name] = t
        return t[:numel]

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _prep(self, t, dtype):
        return t.to(device=self.device, dtype=dtype, non_blocking=True).contiguous()

    # ------------------------------------------------------------------ model/model.py:260-320
    def forward(self, phoneme, melspec, phoneme_lens, melspec_lens, return_alignments: bool = True):
        if self.training:
            raise RuntimeError("transformertts_b200 accelerates eval-mode forward only (dropout off, BatchNorm running "
                               "statistics); call .eval() first. Training-mode forward/backward is out of scope.")
        lib = _ext.load()
        # the target mel may arrive as bf16 (data.collate_fn(mel_dtype=bfloat16)): half the H2D bytes, no cast kernel
        mel_bf16 = melspec.dtype == torch.bfloat16
        if phoneme.dim() != 2 or melspec.dim() != 3 or melspec.shape[2] != N_MELS:
            raise ValueError("expected phoneme [B,Tp] and melspec [B,Tm,80]")
        B, Tp = phoneme.shape
        Tm = melspec.shape[1]
        if melspec.shape[0] != B or phoneme_lens.shape != (B,) or melspec_lens.shape != (B,):
            raise ValueError("batch sizes of phoneme / melspec / lens disagree")
        if self.strict_lens:   # model/model.py:294 asserts the mask width equals the padded length
            assert int(phoneme_lens.max()) == Tp and int(melspec_lens.max()) == Tm, "Size mismatch between memory and mask"
        flags = (FLAG_ENC_MASK | (FLAG_WANT_ALIGN if return_alignments else 0) |
                 (FLAG_SERIAL if self.serial_encoder else 0) | (FLAG_MEL_BF16 if mel_bf16 else 0))
        with torch.cuda.device(self.device):
            if self.use_cuda_graph:                    # inputs are copied (cast / moved) straight into the graph's own buffers
                return self._forward_graph(lib, (phoneme, melspec, phoneme_lens, melspec_lens), B, Tp, Tm, flags,
                                           return_alignments)
            args = (self._prep(phoneme, torch.int64), self._prep(melspec, torch.bfloat16 if mel_bf16 else torch.float32),
                    self._prep(phoneme_lens, torch.int64), self._prep(melspec_lens, torch.int64))
            return self._forward_launch(lib, args, B, Tp, Tm, flags, return_alignments)

    def _forward_launch(self, lib, args, B, Tp, Tm, flags, return_alignments):
        """Enqueue the forward's kernels on the current stream (outputs freshly allocated)."""
        phoneme, melspec, phoneme_lens, melspec_lens = args
        w = self._weights()
        ws, ws_bytes = self._workspace(B, Tp, Tm)
        # pred | post | stop back to back in ONE allocation: a consumer on the host fetches them with one copy
        n = B * Tm * N_MELS
        flat = torch.empty(2 * n + B * Tm, dtype=torch.float32, device=self.device)
        pred, post, stop = flat[:n].view(B, Tm, N_MELS), flat[n:2 * n].view(B, Tm, N_MELS), flat[2 * n:].view(B, Tm)
        align = torch.empty(3, B, 4, Tm, Tp, dtype=torch.float32, device=self.device) if return_alignments else None
        _ext.check(lib.tts_forward(C.byref(w.struct), _ext.ptr(phoneme), _ext.ptr(melspec), _ext.ptr(phoneme_lens),
                                   _ext.ptr(melspec_lens), B, Tp, Tm, flags, _ext.ptr(pred), _ext.ptr(post),
                                   _ext.ptr(stop), _ext.ptr(align), ws, ws_bytes, self._stream()), "tts_forward")
        return {
            "pred_melspec": pred,
            "post_melspec": post,
            "pred_stop": stop,
            "alignments": [align[l] for l in range(3)] if return_alignments else [],
        }

    def _graph_entry(self, lib, B, Tp, Tm, flags, return_alignments) -> "_GraphEntry":
        w = self._weights()
        ws, ws_bytes = self._workspace(B, Tp, Tm)
        key = (B, Tp, Tm, flags, id(w), ws.value)            # shapes and flags only: no caller addresses
        e = self._graphs.get(key)
        if e is not None:
            return e
        if len(self._graphs) >= 8:                           # shapes keep changing (unbucketed loader): do not hoard graphs
            old = self._graphs.pop(next(iter(self._graphs)))
            for _, o in old.sets:
                _STATIC_OUTPUT_READERS.pop((o["pred_melspec"].device.index, o["pred_melspec"].data_ptr()), None)
        dev = self.device
        e = _GraphEntry()
        e.inputs = (torch.zeros(B, Tp, dtype=torch.int64, device=dev),
                    torch.zeros(B, Tm, N_MELS, dtype=torch.bfloat16 if flags & FLAG_MEL_BF16 else torch.float32, device=dev),
                    torch.ones(B, dtype=torch.int64, device=dev), torch.ones(B, dtype=torch.int64, device=dev))
        e.sets, e.turn, e.weights = [], 0, w                 # keep the weights of the capture alive
        self._forward_launch(lib, e.inputs, B, Tp, Tm, flags, return_alignments)    # warm-up outside the capture
        torch.cuda.current_stream(dev).synchronize()
        for _ in range(max(1, int(self.graph_output_sets))):
            g = torch.cuda.CUDAGraph()
            n0 = lib.tts_launch_count()
            with torch.cuda.graph(g):
                out = self._forward_launch(lib, e.inputs, B, Tp, Tm, flags, return_alignments)
            e.launches = int(lib.tts_launch_count() - n0)
            e.sets.append((g, out))
            _STATIC_OUTPUT_READERS[(dev.index, out["pred_melspec"].data_ptr())] = []
        self._graphs[key] = e
        return e

    def graph_input_buffers(self, B, Tp, Tm, mel_dtype=torch.float32, return_alignments: bool = True):
        """The static input tensors (phoneme, melspec, phoneme_lens, melspec_lens) of the graph for these shapes.  A
        producer that writes its batch into them (and passes them to `forward`) saves the per-call input copies."""
        flags = (FLAG_ENC_MASK | (FLAG_WANT_ALIGN if return_alignments else 0) |
                 (FLAG_SERIAL if self.serial_encoder else 0) | (FLAG_MEL_BF16 if mel_dtype == torch.bfloat16 else 0))
        with torch.cuda.device(self.device):
            return self._graph_entry(_ext.load(), B, Tp, Tm, flags, return_alignments).inputs

    def _forward_graph(self, lib, args, B, Tp, Tm, flags, return_alignments):
        """`use_cuda_graph = True`: the ~50 launches of one forward (both streams, programmatic-dependent-launch edges
        included) are captured once per (shapes, flags) and replayed -- 3-6 % less device time, most for small batches.
        The graph reads its OWN input buffers: the caller's tensors (any device, pinned or not, any compatible dtype) are
        copied into them on the current stream before every replay, so fresh tensors from a DataLoader never re-capture
        (`graph_input_buffers` hands the buffers out to producers that want to skip the copy).  The OUTPUTS are static
        buffers too; an entry rotates through `graph_output_sets` (2) of them, so a result stays intact during the next
        call and is overwritten by the one after.  A reader on another stream registers its completion event with
        `note_output_read` (data.OutputFetcher does) and the replay that reuses the set waits for it."""
        e = self._graph_entry(lib, B, Tp, Tm, flags, return_alignments)
        for dst, src in zip(e.inputs, args):
            if src.data_ptr() != dst.data_ptr() or src.device != dst.device:
                dst.copy_(src, non_blocking=True)
        g, out = e.sets[e.turn]
        e.turn = (e.turn + 1) % len(e.sets)
        readers = _STATIC_OUTPUT_READERS.get((self.device.index, out["pred_melspec"].data_ptr()))
        if readers:
            cur = torch.cuda.current_stream(self.device)
            for ev in readers:
                cur.wait_event(ev)
            readers.clear()
        g.replay()
        self.graph_launches += e.launches                    # kernels of this library executed through graph replays
        return {"pred_melspec": out["pred_melspec"], "post_melspec": out["post_melspec"], "pred_stop": out["pred_stop"],
                "alignments": list(out["alignments"])}

    # ------------------------------------------------------------------ model/model.py:323-394
    @torch.no_grad()
    def inference(self, phoneme, phoneme_lens, max_len: int = 1500, stop_threshold: float = 0.5,
                  rows_per_cluster: int = 0, cluster_size: int = 0, forced_melspec=None):
        """`forced_melspec` [B, >= max_len-1, 80] (not in the reference): teacher-forced decode -- step t consumes frame
        t-1 of the given trajectory instead of the model's own prediction, through the same kernel and K/V cache."""
        self.eval()                                   # model/model.py:342
        lib = _ext.load()
        phoneme, phoneme_lens = self._prep(phoneme, torch.int64), self._prep(phoneme_lens, torch.int64)
        if phoneme.dim() != 2 or phoneme_lens.shape != (phoneme.shape[0],):
            raise ValueError("expected phoneme [B,Tp] and phoneme_lens [B]")
        B, Tp = phoneme.shape
        max_frames = int(max_len) - 1                 # `for t in range(1, max_len)`  model/model.py:354
        if max_frames < 1:
            raise ValueError("max_len must be >= 2 (the reference produces at most max_len-1 frames)")
        dev = self.device
        with torch.cuda.device(dev):
            if self._decode_capacity is None:
                self._decode_capacity = int(lib.tts_decode_capacity())
            if B > self._decode_capacity and not self._warned_capacity:
                import warnings
                warnings.warn(f"inference(): batch {B} exceeds the {self._decode_capacity} utterances one decode launch holds "
                              "co-resident; the result is still exact but the launch cannot stop early. Split the batch "
                              "(transformertts_b200.sharding.micro_batches).")
                self._warned_capacity = True
            w = self._weights()
            ws, ws_bytes = self._workspace(B, Tp, 1)
            # decode scratch (K/V caches, raw frame / stop buffers, counters) is kept by the module and only grows: a
            # 128-row, 1000-frame call needs 393 MB of it, which is not re-requested from the allocator on every call
            cross_kv = self._scratch("cross_kv", 3 * B * Tp * 512, torch.bfloat16)
            ph_len = self._scratch("ph_len", B, torch.int32)
            # encoder WITHOUT key-padding mask (model/model.py:346-348)
            _ext.check(lib.tts_encode(C.byref(w.struct), _ext.ptr(phoneme), _ext.ptr(phoneme_lens), B, Tp, 0,
                                      _ext.ptr(cross_kv), _ext.ptr(ph_len), ws, ws_bytes, self._stream()), "tts_encode")
            self_kv = self._scratch("self_kv", 3 * B * max_frames * 512, torch.bfloat16)
            mel = self._scratch("mel", B * max_frames * N_MELS, torch.float32).view(B, max_frames, N_MELS)
            stop = self._scratch("stop", B * max_frames, torch.float32).view(B, max_frames)
            stop_count = self._scratch("stop_count", max_frames, torch.int32)
            n_frames = self._scratch("n_frames", 1, torch.int32)
            forced = None
            if forced_melspec is not None:
                forced = self._prep(forced_melspec, torch.float32)
                if forced.dim() != 3 or forced.shape[0] != B or forced.shape[1] < max_frames or forced.shape[2] != N_MELS:
                    raise ValueError("forced_melspec must be [B, >= max_len-1, 80]")
                forced = forced[:, :max_frames].contiguous()
            _ext.check(lib.tts_decode(C.byref(w.struct), _ext.ptr(cross_kv), _ext.ptr(ph_len), B, Tp, max_frames,
                                      float(stop_threshold), _ext.ptr(self_kv), _ext.ptr(mel), _ext.ptr(stop),
                                      _ext.ptr(stop_count), _ext.ptr(n_frames), int(rows_per_cluster),
                                      int(cluster_size), _ext.ptr(forced), self._stream()), "tts_decode")
            n = int(n_frames.item())                  # the single host sync of the AR loop
            pred = mel[:, :n].clone(memory_format=torch.contiguous_format)       # always a copy: `mel` is module scratch
            pred_stop = stop[:, :n].unsqueeze(-1).clone(memory_format=torch.contiguous_format)   # (B,T,1)  model/model.py:378,393
            post = torch.empty_like(pred)
            ws, ws_bytes = self._workspace(B, 1, n)
            _ext.check(lib.tts_postnet(C.byref(w.struct), _ext.ptr(pred), B, n, _ext.ptr(post), ws, ws_bytes,
                                       self._stream()), "tts_postnet")
        return {"pred_melspec": pred, "post_melspec": post, "pred_stop": pred_stop}
