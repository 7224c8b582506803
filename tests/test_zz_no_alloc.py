"""tts_loss_static: the loss entry point for hosts that must not allocate device memory.

Who needs it: code that embeds libtts_b200.so somewhere it does not own the GPU's memory -- a plugin inside a framework
whose allocator holds (and accounts for) all of HBM, a serving thread that cannot afford the device-wide stall of a
cudaMalloc, a job that has already given every free byte to a caching pool.  Such a caller can still hand over its
input and output buffers, but it has nowhere to get `tts_loss`'s scratch from.  `tts_loss_static` keeps that scratch in
a fixed array inside the library instead.

What is checked, through the C ABI only: the new entry point returns exactly the four scalars the existing `tts_loss`
returns (same kernels, same reduction order, so equality is bit for bit -- tighter than the 2e-5 the loss tests allow
against torch), on two calls in a row with different inputs, because the array is shared and still holds the first
call's partial sums when the second one starts.  The file is named to be collected last: under `pytest -x` a failure
here cannot hide another test.  It passes in a fresh process (its first call does the per-device set-up) and after the
rest of the suite.
"""
import ctypes as C

import pytest
import torch

from transformertts_b200 import _ext

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _case(B, T, seed):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    mel = torch.randn(B, T, 80, generator=g)
    pred = mel + 0.5 * torch.randn(B, T, 80, generator=g)
    post = mel + 0.25 * torch.randn(B, T, 80, generator=g)
    stop = 4 * torch.randn(B, T, generator=g)
    return [t.to(DEV).contiguous() for t in (pred, post, stop, mel, lens)]


def _both(lib, case, B, T, w, stream):
    pred, post, stop, mel, lens = case
    need = lib.tts_loss_scratch_bytes()
    scratch = torch.empty((need + 7) // 8, dtype=torch.float64, device=DEV)       # the existing path: caller scratch
    ref = torch.full((4,), float("nan"), device=DEV)
    got = torch.full((4,), float("nan"), device=DEV)
    args = [_ext.ptr(t) for t in (pred, post, stop, mel, lens)]
    _ext.check(lib.tts_loss(*args, B, T, w, _ext.ptr(ref), _ext.ptr(scratch), scratch.numel() * 8, stream), "tts_loss")
    _ext.check(lib.tts_loss_static(*args, B, T, w, _ext.ptr(got), stream), "tts_loss_static")
    torch.cuda.synchronize()
    return ref.cpu(), got.cpu()


def test_static_scratch_loss_equals_caller_scratch_loss_twice_in_a_row():
    lib = _ext.load()
    with torch.cuda.device(DEV):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        # large first (fills every slot of the shared array), then a tiny batch that most blocks contribute zero to:
        # stale partial sums from call 1 would show up in call 2
        for B, T, w, seed in [(64, 800, 8.0, 1), (3, 129, 4.0, 2)]:
            ref, got = _both(lib, _case(B, T, seed), B, T, w, stream)
            assert torch.isfinite(ref).all() and torch.isfinite(got).all()
            assert torch.equal(ref, got), (B, T, ref.tolist(), got.tolist())


def test_static_scratch_loss_rejects_bad_arguments():
    lib = _ext.load()
    with torch.cuda.device(DEV):
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        pred, post, stop, mel, lens = _case(2, 8, 3)
        out = torch.zeros(4, device=DEV)
        args = [_ext.ptr(t) for t in (pred, post, stop, mel, lens)]
        assert lib.tts_loss_static(*args, 0, 8, 8.0, _ext.ptr(out), stream) != 0
        assert b"positive" in lib.tts_last_error()
        assert lib.tts_loss_static(*args, 2, 8, 8.0, None, stream) != 0
        assert b"NULL" in lib.tts_last_error()
