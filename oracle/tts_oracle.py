"""CPU oracle for the TransformerTTS hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT.

A plain-torch fp32 restatement (functional ops only, no nn.Module) of the
reference's teacher-forced ``TransformerTTS.forward`` and greedy
``TransformerTTS.inference`` (reference: model/model.py:260-320, 323-394).
Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module; the product
package ``transformertts_b200`` never does.

Parity pin: the reference ships no tests or golden vectors,
so this restatement is pinned against the reference module itself, imported
from /root/reference in the build container by ``oracle/make_golden.py``; the
resulting vectors are committed under ``tests/golden/`` and
``tests/test_oracle.py`` re-checks this file against them on every run.

Every function cites the reference lines it follows.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# config.yaml:25-42 (+ ctor defaults model/model.py:157-159)
CONFIG = dict(
    encoder_prenet_n_layers=3, encoder_prenet_in_channel=256,
    encoder_prenet_out_channel=256, encoder_prenet_kernel_size=5,
    encoder_prenet_dropout=0.5, encoder_n_layers=3, encoder_n_head=4,
    encoder_d_ffn=1024, encoder_dropout=0.1, decoder_n_layers=3,
    decoder_n_head=4, decoder_d_ffn=1024, decoder_dropout=0.1,
    postnet_n_layers=5, postnet_kernel_size=5, postnet_dropout=0.5,
    d_model=256,
)
N_PHON, N_MELS, D, H, FFN, KS = 100, 80, 256, 4, 1024, 5
LN_EPS = 1e-5   # torch/nn/modules/transformer.py:750,1037
BN_EPS = 1e-5   # torch/nn/modules/batchnorm.py:41
MAX_PE = 5000   # model/model.py:76


# --------------------------------------------------------------------------
# weights
# --------------------------------------------------------------------------
def pe_table(max_len: int = MAX_PE, d_model: int = D) -> Tensor:
    """model/model.py:80-86."""
    position = torch.arange(0, max_len, dtype=torch.float32).unsqueeze(1)
    div_term = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32)
                         * (-math.log(10000.0)) / d_model)
    pe = torch.zeros(max_len, d_model, dtype=torch.float32)
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe


def _xavier(gen, shape, gain):
    # nn.init.xavier_normal_ (model/module.py:21-22,47-48): std = gain*sqrt(2/(fan_in+fan_out))
    rf = 1
    for s in shape[2:]:
        rf *= s
    fan_in, fan_out = shape[1] * rf, shape[0] * rf
    std = gain * math.sqrt(2.0 / (fan_in + fan_out))
    return torch.randn(shape, generator=gen) * std


def bf16_round(t: Tensor) -> Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def make_state_dict(seed: int = 1234, randomize_bn: bool = True, alpha: float = 0.7,
                    round_bf16: bool = True, stop_bias: float = 0.0,
                    n_enc: int = 3, n_dec: int = 3) -> Dict[str, Tensor]:
    """A state_dict with the reference's key names/shapes,
    drawn from a seeded CPU generator with the reference's init laws
    (model/module.py:21-25,47-50; torch defaults elsewhere).  BN statistics and
    affine terms are randomised so that a wrong BN fold is visible.
    Matrix weights and the embedding are rounded to bf16 (held in fp32) so the
    comparison with the bf16 CUDA path isolates activation rounding.
    """
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    relu_gain, tanh_gain = math.sqrt(2.0), 5.0 / 3.0

    def rnd(t):
        return bf16_round(t) if round_bf16 else t

    def conv_bn(prefix, cin, cout, gain):
        sd[f"{prefix}.conv.weight"] = rnd(_xavier(g, (cout, cin, KS), gain))
        sd[f"{prefix}.conv.bias"] = (torch.randn(cout, generator=g) * 0.05 if randomize_bn
                                     else torch.zeros(cout))
        if randomize_bn:
            sd[f"{prefix}.bn.weight"] = 1.0 + 0.2 * torch.randn(cout, generator=g)
            sd[f"{prefix}.bn.bias"] = 0.1 * torch.randn(cout, generator=g)
            sd[f"{prefix}.bn.running_mean"] = 0.1 * torch.randn(cout, generator=g)
            sd[f"{prefix}.bn.running_var"] = 0.5 + torch.rand(cout, generator=g)
        else:
            sd[f"{prefix}.bn.weight"] = torch.ones(cout)
            sd[f"{prefix}.bn.bias"] = torch.zeros(cout)
            sd[f"{prefix}.bn.running_mean"] = torch.zeros(cout)
            sd[f"{prefix}.bn.running_var"] = torch.ones(cout)
        sd[f"{prefix}.bn.num_batches_tracked"] = torch.zeros((), dtype=torch.int64)

    def linear(prefix, cin, cout, gain=relu_gain, bias_std=0.02):
        sd[f"{prefix}.weight"] = rnd(_xavier(g, (cout, cin), gain))
        sd[f"{prefix}.bias"] = torch.randn(cout, generator=g) * bias_std

    def mha(prefix):
        # nn.MultiheadAttention defaults: xavier_uniform in_proj, zeros bias; biases are
        # randomised here so a dropped bias is visible.
        bound = math.sqrt(6.0 / (3 * D + D))
        sd[f"{prefix}.in_proj_weight"] = rnd((torch.rand(3 * D, D, generator=g) * 2 - 1) * bound)
        sd[f"{prefix}.in_proj_bias"] = torch.randn(3 * D, generator=g) * 0.02
        b2 = 1.0 / math.sqrt(D)
        sd[f"{prefix}.out_proj.weight"] = rnd((torch.rand(D, D, generator=g) * 2 - 1) * b2)
        sd[f"{prefix}.out_proj.bias"] = torch.randn(D, generator=g) * 0.02

    def ffn_ln(prefix, n_norm):
        b1, b2 = 1.0 / math.sqrt(D), 1.0 / math.sqrt(FFN)
        sd[f"{prefix}.linear1.weight"] = rnd((torch.rand(FFN, D, generator=g) * 2 - 1) * b1)
        sd[f"{prefix}.linear1.bias"] = (torch.rand(FFN, generator=g) * 2 - 1) * b1
        sd[f"{prefix}.linear2.weight"] = rnd((torch.rand(D, FFN, generator=g) * 2 - 1) * b2)
        sd[f"{prefix}.linear2.bias"] = (torch.rand(D, generator=g) * 2 - 1) * b2
        for i in range(1, n_norm + 1):
            sd[f"{prefix}.norm{i}.weight"] = 1.0 + 0.1 * torch.randn(D, generator=g)
            sd[f"{prefix}.norm{i}.bias"] = 0.05 * torch.randn(D, generator=g)

    sd["emb.weight"] = rnd(torch.randn(N_PHON, D, generator=g))
    for i in range(3):
        conv_bn(f"enc_prenet.layers.{2 * i}", D, D, relu_gain)
    linear("enc_prenet.linear.linear", D, D)
    linear("dec_prenet.linear1.linear", N_MELS, D)
    linear("dec_prenet.linear2.linear", D, D)
    sd["pe.alpha"] = torch.tensor([alpha], dtype=torch.float32)
    sd["pe.pe"] = pe_table()
    for l in range(n_enc):
        mha(f"encoder.layers.{l}.self_attn")
        ffn_ln(f"encoder.layers.{l}", 2)
    for l in range(n_dec):
        mha(f"decoder.layers.{l}.self_attn")
        mha(f"decoder.layers.{l}.multihead_attn")
        ffn_ln(f"decoder.layers.{l}", 3)
    chans = [N_MELS, D, D, D, D, N_MELS]
    for i in range(5):
        conv_bn(f"postnet.layers.{3 * i}", chans[i], chans[i + 1], tanh_gain)
    linear("linear1.linear", D, N_MELS)
    linear("linear2.linear", D, 1)
    sd["linear2.linear.bias"] = sd["linear2.linear.bias"] + stop_bias
    return sd


def state_dict_checksum(sd: Dict[str, Tensor]) -> float:
    """A cheap fingerprint so fixtures can assert the seeded weights reproduce."""
    tot = 0.0
    for k in sorted(sd):
        if sd[k].dtype.is_floating_point and k != "pe.pe":
            tot += float(sd[k].double().abs().sum())
    return tot


def make_inputs(B: int, Tp: int, Tm: int, seed: int = 1, ragged: bool = True,
                min_frac_p: float = 2 / 3, min_frac_m: float = 5 / 8):
    """Synthetic LJSpeech-shaped batch: phoneme ids in [1,100), zero padded;
    mel ~ N(0,1) rounded to bf16, zero beyond its length; rows sorted by phoneme
    length descending (dataset.py:74-76); row 0 has full Tp and some row full Tm."""
    g = torch.Generator().manual_seed(seed)
    if ragged and B > 1:
        pl = torch.randint(max(1, int(Tp * min_frac_p)), Tp + 1, (B,), generator=g)
        ml = torch.randint(max(1, int(Tm * min_frac_m)), Tm + 1, (B,), generator=g)
        pl[0] = Tp
        ml[int(torch.randint(0, B, (1,), generator=g))] = Tm
        pl, _ = torch.sort(pl, descending=True)
    else:
        pl = torch.full((B,), Tp, dtype=torch.int64)
        ml = torch.full((B,), Tm, dtype=torch.int64)
    ph = torch.randint(1, N_PHON, (B, Tp), generator=g)
    mel = bf16_round(torch.randn(B, Tm, N_MELS, generator=g))
    ar_p = torch.arange(Tp).unsqueeze(0)
    ar_m = torch.arange(Tm).unsqueeze(0)
    ph = ph * (ar_p < pl.unsqueeze(1))
    mel = mel * (ar_m < ml.unsqueeze(1)).unsqueeze(-1)
    return ph.long(), mel.float(), pl.long(), ml.long()


# --------------------------------------------------------------------------
# blocks
# --------------------------------------------------------------------------
def conv_bn(sd, prefix: str, x: Tensor) -> Tensor:
    """ConvNormBN.forward, eval mode (model/module.py:28-33): permute, Conv1d(pad=(K-1)//2),
    BatchNorm1d with running statistics, permute back.  x: (B,T,Cin)."""
    w = sd[f"{prefix}.conv.weight"]
    y = F.conv1d(x.permute(0, 2, 1), w, sd[f"{prefix}.conv.bias"], padding=(w.shape[2] - 1) // 2)
    mean, var = sd[f"{prefix}.bn.running_mean"], sd[f"{prefix}.bn.running_var"]
    y = (y - mean[None, :, None]) / torch.sqrt(var[None, :, None] + BN_EPS)
    y = y * sd[f"{prefix}.bn.weight"][None, :, None] + sd[f"{prefix}.bn.bias"][None, :, None]
    return y.permute(0, 2, 1)


def enc_prenet(sd, x: Tensor) -> Tensor:
    """EncoderPreNet.forward (model/model.py:38-45): 3x[ConvNormBN, Dropout(off)] -> Linear.
    No activation."""
    for i in range(3):
        x = conv_bn(sd, f"enc_prenet.layers.{2 * i}", x)
    return F.linear(x, sd["enc_prenet.linear.linear.weight"], sd["enc_prenet.linear.linear.bias"])


def dec_prenet(sd, x: Tensor) -> Tensor:
    """DecoderPreNet.forward (model/model.py:61-67), dropout off in eval."""
    x = F.relu(F.linear(x, sd["dec_prenet.linear1.linear.weight"], sd["dec_prenet.linear1.linear.bias"]))
    x = F.relu(F.linear(x, sd["dec_prenet.linear2.linear.weight"], sd["dec_prenet.linear2.linear.bias"]))
    return x


def pos_enc(sd, x: Tensor) -> Tensor:
    """PositionalEncoding.forward (model/model.py:91-97); one shared alpha."""
    return x + sd["pe.alpha"] * sd["pe.pe"][: x.size(1), :].unsqueeze(0)


def mha(sd, prefix: str, q_in: Tensor, kv_in: Tensor, mask: Optional[Tensor]):
    """nn.MultiheadAttention as called by the reference (torch/nn/functional.py:5798,
    6620-6665): packed in-proj [Wq;Wk;Wv], q scaled by 1/sqrt(dh) before QK^T, additive
    -inf mask, softmax over keys, per-head probabilities (B,H,Tq,Tk), out-proj.
    mask: bool, broadcastable to (B,1,Tq,Tk), True = masked."""
    W, b = sd[f"{prefix}.in_proj_weight"], sd[f"{prefix}.in_proj_bias"]
    B, Tq, _ = q_in.shape
    Tk = kv_in.shape[1]
    dh = D // H
    q = F.linear(q_in, W[:D], b[:D]).view(B, Tq, H, dh).transpose(1, 2)
    k = F.linear(kv_in, W[D:2 * D], b[D:2 * D]).view(B, Tk, H, dh).transpose(1, 2)
    v = F.linear(kv_in, W[2 * D:], b[2 * D:]).view(B, Tk, H, dh).transpose(1, 2)
    s = (q * (1.0 / math.sqrt(dh))) @ k.transpose(-1, -2)
    if mask is not None:
        s = s.masked_fill(mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = (p @ v).transpose(1, 2).reshape(B, Tq, D)
    o = F.linear(o, sd[f"{prefix}.out_proj.weight"], sd[f"{prefix}.out_proj.bias"])
    return o, p


def _ln(sd, prefix: str, x: Tensor) -> Tensor:
    return F.layer_norm(x, (D,), sd[f"{prefix}.weight"], sd[f"{prefix}.bias"], LN_EPS)


def _ffn(sd, prefix: str, x: Tensor) -> Tensor:
    """_ff_block (torch/nn/modules/transformer.py:980-982)."""
    h = F.relu(F.linear(x, sd[f"{prefix}.linear1.weight"], sd[f"{prefix}.linear1.bias"]))
    return F.linear(h, sd[f"{prefix}.linear2.weight"], sd[f"{prefix}.linear2.bias"])


def encoder(sd, x: Tensor, src_kpm: Optional[Tensor], n_layers: int = 3) -> Tensor:
    """nn.TransformerEncoder, post-norm slow path (torch/nn/modules/transformer.py:951-957)."""
    mask = None if src_kpm is None else src_kpm[:, None, None, :]
    for l in range(n_layers):
        p = f"encoder.layers.{l}"
        a, _ = mha(sd, f"{p}.self_attn", x, x, mask)
        x = _ln(sd, f"{p}.norm1", x + a)
        x = _ln(sd, f"{p}.norm2", x + _ffn(sd, p, x))
    return x


def decoder(sd, x: Tensor, memory: Tensor, tgt_mask: Tensor, tgt_kpm: Tensor,
            src_kpm: Tensor, n_layers: int = 3):
    """TransformerDecoder.forward + TransformerDecoderLayer.forward, post-norm branch
    (model/layers.py:46-50, 85-110); cross-attention returns per-head probabilities
    (model/layers.py:68-74)."""
    self_mask = tgt_mask[None, None, :, :] | tgt_kpm[:, None, None, :]
    cross_mask = src_kpm[:, None, None, :]
    aligns: List[Tensor] = []
    for l in range(n_layers):
        p = f"decoder.layers.{l}"
        a, _ = mha(sd, f"{p}.self_attn", x, x, self_mask)
        x = _ln(sd, f"{p}.norm1", x + a)
        c, A = mha(sd, f"{p}.multihead_attn", x, memory, cross_mask)
        x = _ln(sd, f"{p}.norm2", x + c)
        x = _ln(sd, f"{p}.norm3", x + _ffn(sd, p, x))
        aligns.append(A)
    return x, aligns


def postnet(sd, x: Tensor) -> Tensor:
    """PostNet.forward (model/model.py:129-135): 4x tanh(BN(conv)), then BN(conv)."""
    for i in range(5):
        x = conv_bn(sd, f"postnet.layers.{3 * i}", x)
        if i < 4:
            x = torch.tanh(x)
    return x


def get_mask(phoneme_lens: Optional[Tensor], mel_lens: Optional[Tensor]):
    """TransformerTTS._get_mask (model/model.py:229-257)."""
    src_kpm = tgt_kpm = tgt_mask = None
    if phoneme_lens is not None:
        mp = int(phoneme_lens.max())
        src_kpm = torch.arange(mp).unsqueeze(0).expand(phoneme_lens.size(0), mp) >= phoneme_lens.unsqueeze(1)
    if mel_lens is not None:
        mm = int(mel_lens.max())
        tgt_kpm = torch.arange(mm).unsqueeze(0).expand(mel_lens.size(0), mm) >= mel_lens.unsqueeze(1)
        tgt_mask = torch.triu(torch.ones(mm, mm), diagonal=1).bool()
    return src_kpm, tgt_kpm, tgt_mask


def _n_layers(sd, stack: str) -> int:
    n = 0
    while f"{stack}.layers.{n}.linear1.weight" in sd:
        n += 1
    return n


# --------------------------------------------------------------------------
# entry points
# --------------------------------------------------------------------------
@torch.no_grad()
def forward(sd, phoneme: Tensor, melspec: Tensor, phoneme_lens: Tensor, melspec_lens: Tensor,
            return_intermediates: bool = False) -> dict:
    """TransformerTTS.forward (model/model.py:260-320)."""
    go = torch.zeros_like(melspec[:, :1, :])
    tgt_in = torch.cat((go, melspec[:, :-1, :]), dim=1)                    # :278-279
    src_kpm, tgt_kpm, tgt_mask = get_mask(phoneme_lens, melspec_lens)      # :281-285
    src_in = pos_enc(sd, enc_prenet(sd, sd["emb.weight"][phoneme]))        # :288
    memory = encoder(sd, src_in, src_kpm, _n_layers(sd, "encoder"))        # :289-292
    tgt = pos_enc(sd, dec_prenet(sd, tgt_in))                              # :297
    tgt_out, aligns = decoder(sd, tgt, memory, tgt_mask, tgt_kpm, src_kpm,
                              _n_layers(sd, "decoder"))                    # :298-306
    pred = F.linear(tgt_out, sd["linear1.linear.weight"], sd["linear1.linear.bias"])  # :309
    post = postnet(sd, pred) + pred                                        # :310
    stop = F.linear(tgt_out, sd["linear2.linear.weight"], sd["linear2.linear.bias"]).squeeze(-1)  # :313
    out = {"pred_melspec": pred, "post_melspec": post, "pred_stop": stop, "alignments": aligns}
    if return_intermediates:
        out.update(src_in=src_in, memory=memory, tgt_in=tgt, tgt_out=tgt_out)
    return out


@torch.no_grad()
def inference(sd, phoneme: Tensor, phoneme_lens: Tensor, max_len: int = 1500,
              stop_threshold: float = 0.5) -> dict:
    """TransformerTTS.inference (model/model.py:323-394), restated faithfully: encoder
    WITHOUT key-padding mask, whole prefix recomputed every step (no KV cache,
    F7), break when all rows' sigmoid(stop) >= threshold at the same step (F9), PostNet once
    at the end, pred_stop returned as (B,T,1)."""
    B = phoneme.size(0)
    n_dec = _n_layers(sd, "decoder")
    src_in = pos_enc(sd, enc_prenet(sd, sd["emb.weight"][phoneme]))        # :345
    memory = encoder(sd, src_in, None, _n_layers(sd, "encoder"))           # :346-348
    ys = [torch.zeros(B, 1, N_MELS)]                                       # :351
    stops = []
    for t in range(1, max_len):                                            # :354
        tgt = pos_enc(sd, dec_prenet(sd, torch.cat(ys, dim=1)))            # :355-356
        src_kpm, tgt_kpm, tgt_mask = get_mask(phoneme_lens, torch.tensor([t] * B))
        tgt_out, _ = decoder(sd, tgt, memory, tgt_mask, tgt_kpm, src_kpm, n_dec)
        cur = tgt_out[:, -1:, :]                                           # :376
        pred_mel = F.linear(cur, sd["linear1.linear.weight"], sd["linear1.linear.bias"])
        pred_stop = F.linear(cur, sd["linear2.linear.weight"], sd["linear2.linear.bias"]).squeeze(-1)
        ys.append(pred_mel)
        stops.append(pred_stop)
        if bool((torch.sigmoid(pred_stop) >= stop_threshold).all()):       # :382-384
            break
    pred = torch.cat(ys[1:], dim=1)                                        # :387
    post = postnet(sd, pred) + pred                                        # :388
    return {"pred_melspec": pred, "post_melspec": post, "pred_stop": torch.stack(stops, dim=1)}


@torch.no_grad()
def inference_cached(sd, phoneme: Tensor, phoneme_lens: Tensor, max_len: int = 1500,
                     stop_threshold: float = 0.5) -> dict:
    """Same result as :func:`inference` (the decoder is causal and eval-mode deterministic,
    SURVEY F7) computed with a per-layer K/V cache, so that long AR parity cases finish in
    seconds on CPU.  tests/test_oracle.py checks it against :func:`inference`."""
    B, Tp = phoneme.shape
    n_dec = _n_layers(sd, "decoder")
    dh = D // H
    src_in = pos_enc(sd, enc_prenet(sd, sd["emb.weight"][phoneme]))
    memory = encoder(sd, src_in, None, _n_layers(sd, "encoder"))
    src_kpm = (torch.arange(Tp).unsqueeze(0) >= phoneme_lens.unsqueeze(1))[:, None, None, :]
    ck, cv, sk, sv = [], [], [], []
    for l in range(n_dec):
        W, b = (sd[f"decoder.layers.{l}.multihead_attn.in_proj_weight"],
                sd[f"decoder.layers.{l}.multihead_attn.in_proj_bias"])
        ck.append(F.linear(memory, W[D:2 * D], b[D:2 * D]).view(B, Tp, H, dh).transpose(1, 2))
        cv.append(F.linear(memory, W[2 * D:], b[2 * D:]).view(B, Tp, H, dh).transpose(1, 2))
        sk.append(torch.zeros(B, H, 0, dh))
        sv.append(torch.zeros(B, H, 0, dh))
    pe, alpha = sd["pe.pe"], sd["pe.alpha"]
    prev = torch.zeros(B, 1, N_MELS)
    mels, stops = [], []
    scale = 1.0 / math.sqrt(dh)
    for t in range(1, max_len):
        x = dec_prenet(sd, prev) + alpha * pe[t - 1][None, None, :]
        for l in range(n_dec):
            p = f"decoder.layers.{l}"
            W, b = sd[f"{p}.self_attn.in_proj_weight"], sd[f"{p}.self_attn.in_proj_bias"]
            qkv = F.linear(x, W, b)
            q, k, v = [z.view(B, 1, H, dh).transpose(1, 2) for z in qkv.split(D, dim=-1)]
            sk[l] = torch.cat((sk[l], k), dim=2)
            sv[l] = torch.cat((sv[l], v), dim=2)
            a = torch.softmax((q * scale) @ sk[l].transpose(-1, -2), dim=-1) @ sv[l]
            a = F.linear(a.transpose(1, 2).reshape(B, 1, D), sd[f"{p}.self_attn.out_proj.weight"],
                         sd[f"{p}.self_attn.out_proj.bias"])
            x = _ln(sd, f"{p}.norm1", x + a)
            W, b = sd[f"{p}.multihead_attn.in_proj_weight"], sd[f"{p}.multihead_attn.in_proj_bias"]
            q = F.linear(x, W[:D], b[:D]).view(B, 1, H, dh).transpose(1, 2)
            s = ((q * scale) @ ck[l].transpose(-1, -2)).masked_fill(src_kpm, float("-inf"))
            c = (torch.softmax(s, dim=-1) @ cv[l]).transpose(1, 2).reshape(B, 1, D)
            c = F.linear(c, sd[f"{p}.multihead_attn.out_proj.weight"], sd[f"{p}.multihead_attn.out_proj.bias"])
            x = _ln(sd, f"{p}.norm2", x + c)
            x = _ln(sd, f"{p}.norm3", x + _ffn(sd, p, x))
        pred_mel = F.linear(x, sd["linear1.linear.weight"], sd["linear1.linear.bias"])
        pred_stop = F.linear(x, sd["linear2.linear.weight"], sd["linear2.linear.bias"]).squeeze(-1)
        mels.append(pred_mel)
        stops.append(pred_stop)
        prev = pred_mel
        if bool((torch.sigmoid(pred_stop) >= stop_threshold).all()):
            break
    pred = torch.cat(mels, dim=1)
    post = postnet(sd, pred) + pred
    return {"pred_melspec": pred, "post_melspec": post, "pred_stop": torch.stack(stops, dim=1)}


def plant_stop_bias(sd, phoneme: Tensor, phoneme_lens: Tensor, horizon: int, min_t: int = 8,
                    delta: float = 0.25):
    """Choose ``linear2.linear.bias`` so the AR loop stops at a known step with a logit
    margin >= delta on both sides (SURVEY 8(c) "AR stop-index parity"; F10: at random init
    the stop logit never crosses for B>1).  The stop logit is not fed back, so shifting the
    bias leaves the trajectory unchanged.  Returns (bias_value, n_frames, margin)."""
    sd2 = dict(sd)
    sd2["linear2.linear.bias"] = torch.tensor([-1.0e4])
    z = inference_cached(sd2, phoneme, phoneme_lens, max_len=horizon + 1)["pred_stop"][..., 0] + 1.0e4
    return plant_from_logits(z, min_t=min_t, delta=delta)


def plant_from_logits(z: Tensor, min_t: int = 8, delta: float = 0.25):
    """The chooser half of :func:`plant_stop_bias`: z (B,T) are the bias-free stop logits of a never-stopping run.
    Picks the step >= min_t at which the all-rows minimum makes its largest new maximum (gap >= 2 delta) and returns
    (bias_value, n_frames, margin) so that step is the first one where every row crosses."""
    m = z.min(dim=0).values                      # (T,) the all-rows rule looks at the min
    best = None
    run_max = float(m[0])
    for t in range(1, m.numel()):
        gap = float(m[t]) - run_max
        if t + 1 >= min_t and gap >= 2 * delta and (best is None or gap > best[1]):
            best = (t, gap, run_max)
        run_max = max(run_max, float(m[t]))
    if best is None:
        raise RuntimeError("no stop step with the requested margin inside the horizon")
    t, gap, prev_max = best
    bias = -(float(m[t]) + prev_max) / 2.0
    return bias, t + 1, gap / 2.0
