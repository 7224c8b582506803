"""Shared helpers of the GPU parity tests: call the CUDA path through the C ABI (ctypes) and compare
with plain fp32 torch on the CPU (per operator) or with the oracle (whole path).  Test infrastructure."""
import ctypes as C
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

from transformertts_b200 import _ext  # noqa: E402

DEV = "cuda:0"


def bf16r(t):
    return t.to(torch.bfloat16).to(torch.float32)


def dev_bf16(t):
    return t.to(torch.bfloat16).contiguous().to(DEV)


def dev_f32(t):
    return t.to(torch.float32).contiguous().to(DEV)


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def report(name, got, ref, atol, rtol=0.0):
    """Returns (ok, message) with the location of the worst element."""
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    if got.shape != ref.shape:
        return False, f"{name}: shape {tuple(got.shape)} != {tuple(ref.shape)}"
    if not torch.isfinite(got).all():
        bad = (~torch.isfinite(got)).sum().item()
        return False, f"{name}: {bad} non-finite values of {got.numel()}"
    err = (got - ref).abs()
    tol = atol + rtol * ref.abs()
    worst = int(torch.argmax(err - tol))
    idx = np.unravel_index(worst, tuple(ref.shape)) if ref.dim() > 0 else ()
    ok = bool((err <= tol).all())
    rel_l2 = float((got - ref).norm() / (ref.norm() + 1e-20))
    return ok, (f"{name}: max_abs_err={float(err.max()):.4g} rel_l2={rel_l2:.3g} ref_absmax={float(ref.abs().max()):.4g} "
                f"worst@{tuple(int(i) for i in idx)} got={float(got.flatten()[worst]):.5g} "
                f"ref={float(ref.flatten()[worst]):.5g} atol={atol} frac_bad={float((err > tol).float().mean()):.3g}")


# ---------------------------------------------------------------- operator wrappers (C ABI)
def op_linear(A, W, bias, act=0, pe=None, alpha=0.0, pe_period=0, residual=None, ln=None, want_f32=False):
    """A [M,K], W [N,K] fp32 cpu tensors (already bf16-representable)."""
    lib = _ext.load()
    M, K = A.shape
    N = W.shape[0]
    Kp = (K + 63) // 64 * 64
    Ad = torch.zeros(M, Kp); Ad[:, :K] = A
    Wd = torch.zeros(N, Kp); Wd[:, :K] = W
    Ad, Wd, bd = dev_bf16(Ad), dev_bf16(Wd), dev_f32(bias)
    ped = dev_f32(pe) if pe is not None else None
    rd = dev_bf16(residual) if residual is not None else None
    g = dev_f32(ln[0]) if ln is not None else None
    b = dev_f32(ln[1]) if ln is not None else None
    # one kernel family, two output plans: fp32 (+ optional bf16 copy) for N <= 128, bf16 only for N % 256 == 0
    o16 = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV) if (not want_f32 or N % 8 == 0) else None
    o32 = torch.full((M, N), float("nan"), dtype=torch.float32, device=DEV) if want_f32 else None
    _ext.check(lib.tts_op_linear(_ext.ptr(Ad), _ext.ptr(Wd), _ext.ptr(bd), M, N, Kp, act, _ext.ptr(ped), float(alpha),
                                 int(pe_period), _ext.ptr(rd), _ext.ptr(g), _ext.ptr(b), _ext.ptr(o16), _ext.ptr(o32),
                                 stream()), "tts_op_linear")
    torch.cuda.synchronize()
    return (o16.float().cpu() if o16 is not None else None), (o32.cpu() if want_f32 else None)


def ref_linear(A, W, bias, act=0, pe=None, alpha=0.0, pe_period=0, residual=None, ln=None):
    y = A.double() @ W.double().t() + bias.double()
    if act == 1:
        y = torch.relu(y)
    elif act == 2:
        y = torch.tanh(y)
    if pe is not None:
        pos = torch.arange(A.shape[0]) % pe_period
        y = y + alpha * pe.double()[pos]
    if residual is not None:
        y = y + residual.double()
    if ln is not None:
        y = torch.nn.functional.layer_norm(y, (y.shape[1],), ln[0].double(), ln[1].double(), 1e-5)
    return y.float()


def op_linear_chain(A, W1, b1, W2, b2, act1=0, residual=None, ln=None, act2=0, pe=None, alpha=0.0, pe_period=0, store1=True):
    """Two chained GEMMs in one launch (tts_op_linear_chain).  Returns (y1 or None, out2) as fp32 cpu tensors."""
    lib = _ext.load()
    M, K = A.shape
    Kp = (K + 63) // 64 * 64
    Ad = torch.zeros(M, Kp); Ad[:, :K] = A
    W1d = torch.zeros(256, Kp); W1d[:, :K] = W1
    Ad, W1d, W2d = dev_bf16(Ad), dev_bf16(W1d), dev_bf16(W2)
    b1d, b2d = dev_f32(b1), dev_f32(b2)
    rd = dev_bf16(residual) if residual is not None else None
    g = dev_f32(ln[0]) if ln is not None else None
    b = dev_f32(ln[1]) if ln is not None else None
    ped = dev_f32(pe) if pe is not None else None
    o1 = torch.full((M, 256), float("nan"), dtype=torch.bfloat16, device=DEV) if store1 else None
    o2 = torch.full((M, 256), float("nan"), dtype=torch.bfloat16, device=DEV)
    _ext.check(lib.tts_op_linear_chain(_ext.ptr(Ad), _ext.ptr(W1d), _ext.ptr(b1d), M, Kp, act1, _ext.ptr(rd), _ext.ptr(g),
                                       _ext.ptr(b), _ext.ptr(W2d), _ext.ptr(b2d), act2, _ext.ptr(ped), float(alpha),
                                       int(pe_period), _ext.ptr(o1), _ext.ptr(o2), stream()), "tts_op_linear_chain")
    torch.cuda.synchronize()
    return (o1.float().cpu() if store1 else None), o2.float().cpu()


def op_conv1d(x, w, bias, act=0, residual=None, want_f32=False):
    """x [B,T,C], w [N,C,taps] (torch Conv1d layout) fp32 cpu (bf16-representable)."""
    lib = _ext.load()
    B, T, Cc = x.shape
    N, _, taps = w.shape
    Cp = (Cc + 63) // 64 * 64
    wp = torch.zeros(N, taps, Cp)
    wp[:, :, :Cc] = w.permute(0, 2, 1)
    xd, wd, bd = dev_bf16(x), dev_bf16(wp.reshape(N, taps * Cp)), dev_f32(bias)
    rd = dev_f32(residual) if residual is not None else None
    o16 = torch.full((B, T, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    o32 = torch.full((B, T, N), float("nan"), dtype=torch.float32, device=DEV) if want_f32 else None
    _ext.check(lib.tts_op_conv1d(_ext.ptr(xd), _ext.ptr(wd), _ext.ptr(bd), B, T, Cc, N, taps, act, _ext.ptr(rd),
                                 _ext.ptr(o16), _ext.ptr(o32), stream()), "tts_op_conv1d")
    torch.cuda.synchronize()
    return o16.float().cpu(), (o32.cpu() if want_f32 else None)


def ref_conv1d(x, w, bias, act=0, residual=None):
    y = torch.nn.functional.conv1d(x.double().permute(0, 2, 1), w.double(), bias.double(),
                                   padding=(w.shape[2] - 1) // 2).permute(0, 2, 1)
    if act == 1:
        y = torch.relu(y)
    elif act == 2:
        y = torch.tanh(y)
    if residual is not None:
        y = y + residual.double()
    return y.float()


def op_attention(q, k, v, klen=None, causal=False, want_probs=False):
    """q [B,Tq,256], k/v [B,Tk,256] fp32 cpu (bf16-representable; q already scaled by 1/8)."""
    lib = _ext.load()
    B, Tq, _ = q.shape
    Tk = k.shape[1]
    qd, kd, vd = dev_bf16(q), dev_bf16(k), dev_bf16(v)
    kl = klen.to(torch.int32).to(DEV) if klen is not None else None
    out = torch.full((B, Tq, 256), float("nan"), dtype=torch.bfloat16, device=DEV)
    probs = torch.full((B, 4, Tq, Tk), float("nan"), dtype=torch.float32, device=DEV) if want_probs else None
    _ext.check(lib.tts_op_attention(_ext.ptr(qd), 256, _ext.ptr(kd), 256, _ext.ptr(vd), 256, B, Tq, Tk, _ext.ptr(kl),
                                    1 if causal else 0, _ext.ptr(out), _ext.ptr(probs), stream()), "tts_op_attention")
    torch.cuda.synchronize()
    return out.float().cpu(), (probs.cpu() if want_probs else None)


def ref_attention(q, k, v, klen=None, causal=False):
    B, Tq, _ = q.shape
    Tk = k.shape[1]
    qh = q.double().view(B, Tq, 4, 64).transpose(1, 2)
    kh = k.double().view(B, Tk, 4, 64).transpose(1, 2)
    vh = v.double().view(B, Tk, 4, 64).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2)
    mask = torch.zeros(B, 1, Tq, Tk, dtype=torch.bool)
    if klen is not None:
        mask = mask | (torch.arange(Tk)[None, :] >= klen[:, None])[:, None, None, :]
    if causal:
        mask = mask | torch.triu(torch.ones(Tq, Tk), diagonal=1).bool()[None, None]
    s = s.masked_fill(mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = (p @ vh).transpose(1, 2).reshape(B, Tq, 256)
    return o.float(), p.float()


def op_ffn(x, w1, b1, w2, b2, ln):
    """Fused FFN block through tts_op_ffn: LayerNorm(x + W2 relu(W1 x + b1) + b2), bf16 out."""
    lib = _ext.load()
    M = x.shape[0]
    xd, w1d, w2d = dev_bf16(x), dev_bf16(w1), dev_bf16(w2)
    b1d, b2d, g, b = dev_f32(b1), dev_f32(b2), dev_f32(ln[0]), dev_f32(ln[1])
    out = torch.full((M, 256), float("nan"), dtype=torch.bfloat16, device=DEV)
    _ext.check(lib.tts_op_ffn(_ext.ptr(xd), _ext.ptr(w1d), _ext.ptr(b1d), _ext.ptr(w2d), _ext.ptr(b2d), _ext.ptr(g),
                              _ext.ptr(b), M, _ext.ptr(out), stream()), "tts_op_ffn")
    torch.cuda.synchronize()
    return out.float().cpu()


def ref_ffn(x, w1, b1, w2, b2, ln, round_hidden=True):
    h = torch.relu(x.double() @ w1.double().t() + b1.double())
    if round_hidden:                      # the kernel keeps the hidden activation in bf16 (A operand of the 2nd GEMM)
        h = h.float().to(torch.bfloat16).double()
    y = x.double() + h @ w2.double().t() + b2.double()
    y = torch.nn.functional.layer_norm(y, (256,), ln[0].double(), ln[1].double(), 1e-5)
    return y.float()
