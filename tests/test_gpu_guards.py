"""Out-of-bounds WRITE detection without compute-sanitizer (closed on this pool: profiles/r02_sanitizer_closed.txt).
Every buffer the C ABI writes -- outputs, alignments, workspace, K/V caches -- is carved out of one arena with 4 KiB
guard bands filled with a sentinel on both sides; after tts_forward / tts_encode / tts_decode / tts_postnet / tts_loss /
tts_teacher_forcing_mix / tts_pack_outputs the bands must be untouched, for ragged shapes that do not divide the
128-row / 64-key tiles (the TMA stores clip at tensor bounds; the bulk stores and the plain stores must do the same)."""
import ctypes as C

import pytest
import torch

import tts_oracle as O
from parity_common import DEV
from transformertts_b200 import _ext
from transformertts_b200.weights import PackedWeights

pytestmark = pytest.mark.gpu
GUARD = 4096
SENT = 0x5A


class Arena:
    def __init__(self, nbytes):
        self.buf = torch.full((nbytes,), SENT, dtype=torch.uint8, device=DEV)
        self.off = 0
        self.regions = []

    def take(self, nbytes, align=1024):
        base = self.buf.data_ptr()
        start = (base + self.off + GUARD + align - 1) // align * align - base
        self.regions.append((start, nbytes))
        self.off = start + nbytes
        assert self.off + GUARD <= self.buf.numel()
        return C.c_void_p(base + start), self.buf[start:start + nbytes]

    def check(self):
        mask = torch.ones(self.buf.numel(), dtype=torch.bool, device=DEV)
        for s, n in self.regions:
            mask[s:s + n] = False
        bad = (self.buf != SENT) & mask
        assert not bool(bad.any()), f"{int(bad.sum())} guard bytes overwritten, first at {int(bad.nonzero()[0])}"


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("B,Tp,Tm", [(3, 37, 131), (2, 150, 257), (1, 5, 1), (5, 190, 300)])
def test_forward_writes_stay_inside_their_buffers(B, Tp, Tm):
    lib = _ext.load()
    w = PackedWeights(O.make_state_dict(7), DEV)
    ph, mel, pl, ml = (t.to(DEV) for t in O.make_inputs(B, Tp, Tm, seed=8))
    ws_bytes = lib.tts_workspace_bytes(B, Tp, Tm)
    n = B * Tm
    A = Arena(ws_bytes + (3 * n * 80 + n + 3 * B * 4 * Tm * Tp) * 4 + n * 168 * 2 + lib.tts_loss_scratch_bytes()
              + 16 * (GUARD + 1024))
    ws, _ = A.take(ws_bytes)
    pre, pre_t = A.take(n * 80 * 4)
    post, post_t = A.take(n * 80 * 4)
    stop, _ = A.take(n * 4)
    al, al_t = A.take(3 * B * 4 * Tm * Tp * 4)
    _ext.check(lib.tts_forward(C.byref(w.struct), _ext.ptr(ph), _ext.ptr(mel), _ext.ptr(pl), _ext.ptr(ml), B, Tp, Tm, 1 | 2,
                               pre, post, stop, al, ws, ws_bytes, _stream()), "tts_forward")
    torch.cuda.synchronize()
    A.check()
    assert torch.isfinite(post_t.view(torch.float32)).all()
    rows = al_t.view(torch.float32).view(3, B, 4, Tm, Tp).sum(-1)
    assert float((rows - 1).abs().max()) <= 1e-3
    # aux kernels on the same outputs
    out4, _ = A.take(16)
    scratch_bytes = lib.tts_loss_scratch_bytes()
    scratch, _ = A.take(scratch_bytes)
    _ext.check(lib.tts_loss(pre, post, stop, _ext.ptr(mel), _ext.ptr(ml), B, Tm, 8.0, out4, scratch, scratch_bytes, _stream()),
               "tts_loss")
    mixed, _ = A.take(n * 80 * 4)
    seed = (torch.rand(B, Tm, device=DEV) < 0.3).to(torch.uint8)
    _ext.check(lib.tts_teacher_forcing_mix(pre, _ext.ptr(mel), _ext.ptr(seed), _ext.ptr(ml), B, Tm, 8, mixed, _stream()),
               "tts_teacher_forcing_mix")
    packed, _ = A.take(n * 168 * 2)
    _ext.check(lib.tts_pack_outputs(pre, post, stop, n, packed, _stream()), "tts_pack_outputs")
    torch.cuda.synchronize()
    A.check()


@pytest.mark.parametrize("B,Tp,frames,R,C_", [(3, 37, 19, 0, 0), (9, 20, 11, 8, 4), (5, 150, 33, 2, 8), (1, 7, 40, 1, 16)])
def test_inference_writes_stay_inside_their_buffers(B, Tp, frames, R, C_):
    lib = _ext.load()
    w = PackedWeights(O.make_state_dict(9, stop_bias=-1e4), DEV)
    ph, _, pl, _ = (t.to(DEV) for t in O.make_inputs(B, Tp, 8, seed=10))
    ws_bytes = max(lib.tts_workspace_bytes(B, Tp, 1), lib.tts_workspace_bytes(B, 1, frames))
    A = Arena(ws_bytes + 3 * B * (Tp + frames) * 512 * 2 + 3 * B * frames * 80 * 4 + 64 * 1024 + 16 * (GUARD + 1024))
    ws, _ = A.take(ws_bytes)
    ckv, _ = A.take(3 * B * Tp * 512 * 2)
    plen, _ = A.take(B * 4)
    skv, _ = A.take(3 * B * frames * 512 * 2)
    mel, mel_t = A.take(B * frames * 80 * 4)
    stop, _ = A.take(B * frames * 4)
    cnt, _ = A.take(frames * 4)
    nfr, nfr_t = A.take(4)
    post, post_t = A.take(B * frames * 80 * 4)
    _ext.check(lib.tts_encode(C.byref(w.struct), _ext.ptr(ph), _ext.ptr(pl), B, Tp, 0, ckv, plen, ws, ws_bytes, _stream()), "tts_encode")
    _ext.check(lib.tts_decode(C.byref(w.struct), ckv, plen, B, Tp, frames, 0.5, skv, mel, stop, cnt, nfr, R, C_, None, _stream()),
               "tts_decode")
    _ext.check(lib.tts_postnet(C.byref(w.struct), mel, B, frames, post, ws, ws_bytes, _stream()), "tts_postnet")
    torch.cuda.synchronize()
    A.check()
    assert int(nfr_t.view(torch.int32)[0]) == frames
    assert torch.isfinite(post_t.view(torch.float32)).all() and torch.isfinite(mel_t.view(torch.float32)).all()
