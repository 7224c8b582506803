"""Device time of every operator shape of one cfg-2 forward (B=64, Tp=150, Tm=800), each timed alone with CUDA
events and an L2 flush between launches.  Usage: python tools/time_ops.py [filter]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformertts_b200 import _ext

lib = _ext.load()
dev = "cuda:0"
B, Tp, Tm = 64, 150, 800
Np, Nm = B * Tp, B * Tm
D, F = 256, 1024
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
P = _ext.ptr


WARM = bool(os.environ.get("TTS_TIME_OPS_WARM"))   # skip the L2 flush: operands of the previous repetition stay cached


def timeit(fn, reps=8):
    for _ in range(2):
        fn()
    tot = 0.0
    for _ in range(reps):
        if not WARM:
            flush_buf.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize()
        tot += a.elapsed_time(b)
    return tot / reps * 1e3   # us


def bf(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


def linear(M, N, K, act=0, ln=False, pe=False, f32=False):
    A, W, bias = bf(M, K), bf(N, K, scale=K ** -0.5), torch.zeros(N, device=dev)
    res = bf(M, N) if ln else None
    g = torch.ones(N, device=dev) if ln else None
    bb = torch.zeros(N, device=dev) if ln else None
    pet = torch.randn(5000, 256, device=dev) if pe else None
    o16 = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    o32 = torch.empty(M, N, device=dev) if f32 else None

    def fn():
        _ext.check(lib.tts_op_linear(P(A), P(W), P(bias), M, N, K, act, P(pet), 0.7, 800 if pe else 0, P(res), P(g), P(bb),
                                     P(o16), P(o32), st()), "linear")
    fn.keep = (A, W, bias, res, g, bb, pet, o16, o32)
    return fn, 2.0 * M * N * K


def chain(M, K1, ln=False, pe=False, store1=True):
    A, W1, b1 = bf(M, K1), bf(256, K1, scale=K1 ** -0.5), torch.zeros(256, device=dev)
    W2, b2 = bf(256, 256, scale=1 / 16), torch.zeros(256, device=dev)
    res = bf(M, 256) if ln else None
    g = torch.ones(256, device=dev) if ln else None
    bb = torch.zeros(256, device=dev) if ln else None
    pet = torch.randn(5000, 256, device=dev) if pe else None
    o1 = torch.empty(M, 256, dtype=torch.bfloat16, device=dev) if store1 else None
    o2 = torch.empty(M, 256, dtype=torch.bfloat16, device=dev)

    def fn():
        _ext.check(lib.tts_op_linear_chain(P(A), P(W1), P(b1), M, K1, 0 if ln else 1, P(res), P(g), P(bb), P(W2), P(b2),
                                           1 if pe else 0, P(pet), 0.7, 800 if pe else 0, P(o1), P(o2), st()), "chain")
    fn.keep = (A, W1, b1, W2, b2, res, g, bb, pet, o1, o2)
    return fn, 2.0 * M * 256 * (K1 + 256)


def ffn(M):
    x, w1, w2 = bf(M, 256), bf(1024, 256, scale=1 / 16), bf(256, 1024, scale=1 / 32)
    b1, b2, g, b = (torch.zeros(1024, device=dev), torch.zeros(256, device=dev), torch.ones(256, device=dev),
                    torch.zeros(256, device=dev))
    out = torch.empty(M, 256, dtype=torch.bfloat16, device=dev)

    def fn():
        _ext.check(lib.tts_op_ffn(P(x), P(w1), P(b1), P(w2), P(b2), P(g), P(b), M, P(out), st()), "ffn")
    fn.keep = (x, w1, w2, b1, b2, g, b, out)
    return fn, 2.0 * M * 2 * 256 * 1024


def conv(Bc, T, Cc, N, act=0, f32=False):
    Cp = (Cc + 63) // 64 * 64
    x, W, bias = bf(Bc, T, Cc), bf(N, 5 * Cp, scale=(5 * Cc) ** -0.5), torch.zeros(N, device=dev)
    o16 = torch.empty(Bc, T, N, dtype=torch.bfloat16, device=dev) if not f32 else None
    o32 = torch.empty(Bc, T, N, device=dev) if f32 else None

    def fn():
        _ext.check(lib.tts_op_conv1d(P(x), P(W), P(bias), Bc, T, Cc, N, 5, act, None, P(o16), P(o32), st()), "conv")
    fn.keep = (x, W, bias, o16, o32)
    return fn, 2.0 * Bc * T * N * 5 * Cc


def attn(Tq, Tk, causal, probs, packed):
    if packed:
        qkv = bf(B, Tq, 768, scale=0.5)
        q, k, v, ld = qkv[:, :, :256], qkv[:, :, 256:512], qkv[:, :, 512:], 768
        keep = (qkv,)
    else:
        qq, kv = bf(B, Tq, 256, scale=0.5), bf(B, Tk, 512, scale=0.5)
        q, k, v, ld = qq, kv[:, :, :256], kv[:, :, 256:], 512
        keep = (qq, kv)
    o = torch.empty(B, Tq, 256, dtype=torch.bfloat16, device=dev)
    pr = torch.empty(B, 4, Tq, Tk, device=dev) if probs else None
    klen = torch.randint(int(Tk * 0.66), Tk + 1, (B,), device=dev).to(torch.int32)
    ldq = 768 if packed else 256

    def fn():
        _ext.check(lib.tts_op_attention(C.c_void_p(q.data_ptr()), ldq, C.c_void_p(k.data_ptr()), ld,
                                        C.c_void_p(v.data_ptr()), ld, B, Tq, Tk, P(klen), causal, P(o), P(pr), st()), "attn")
    fn.keep = keep + (o, pr, klen)
    return fn, 4.0 * B * Tq * Tk * 256 * (0.5 if causal else 1.0)


OPS = [
    ("enc conv 9600 256->256 k5", lambda: conv(B, Tp, 256, 256), 3),
    ("enc linear+PE 9600x256x256", lambda: linear(Np, 256, 256, pe=True), 1),
    ("enc qkv 9600x768x256", lambda: linear(Np, 768, 256), 3),
    ("enc attn 150x150", lambda: attn(Tp, Tp, 0, False, True), 3),
    ("enc out+LN 9600x256x256", lambda: linear(Np, 256, 256, ln=True), 3),
    ("enc ffn fused 9600x256->1024->256 +LN", lambda: ffn(Np), 3),
    ("enc ffn1 9600x1024x256 (unfused, not in the forward)", lambda: linear(Np, 1024, 256, act=1), 0),
    ("enc ffn2+LN 9600x256x1024 (unfused)", lambda: linear(Np, 256, 1024, ln=True), 0),
    ("dec prenet1 51200x256x128", lambda: linear(Nm, 256, 128, act=1), 1),
    ("dec prenet2+PE 51200x256x256", lambda: linear(Nm, 256, 256, act=1, pe=True), 1),
    ("dec qkv 51200x768x256", lambda: linear(Nm, 768, 256), 3),
    ("dec self-attn causal 800x800", lambda: attn(Tm, Tm, 1, False, True), 3),
    ("dec out+LN 51200x256x256", lambda: linear(Nm, 256, 256, ln=True), 6),
    ("dec cq 51200x256x256", lambda: linear(Nm, 256, 256), 3),
    ("dec ckv 9600x512x256", lambda: linear(Np, 512, 256), 3),
    ("chain dec out+LN -> cq 51200 (one launch for the two above)", lambda: chain(Nm, 256, ln=True), 0),
    ("chain dec prenet1 -> prenet2+PE 51200", lambda: chain(Nm, 128, pe=True, store1=False), 0),
    ("dec cross-attn 800x150 +probs", lambda: attn(Tm, Tp, 0, True, False), 3),
    ("dec ffn fused 51200x256->1024->256 +LN", lambda: ffn(Nm), 3),
    ("dec ffn1 51200x1024x256 (unfused, not in the forward)", lambda: linear(Nm, 1024, 256, act=1), 0),
    ("dec ffn2+LN 51200x256x1024 (unfused)", lambda: linear(Nm, 256, 1024, ln=True), 0),
    ("heads 51200x80x256 f32+bf16", lambda: linear(Nm, 80, 256, f32=True), 1),
    ("postnet conv 80->256 tanh", lambda: conv(B, Tm, 80, 256, act=2), 1),
    ("postnet conv 256->256 tanh", lambda: conv(B, Tm, 256, 256, act=2), 3),
    ("postnet conv 256->80 f32", lambda: conv(B, Tm, 256, 80, f32=True), 1),
]

if os.environ.get("TTS_TIME_OPS_EXTRA"):      # per-tile cost of the streaming attention kernel: time vs. key tiles per item
    OPS += [("x self-attn dense 800x800", lambda: attn(Tm, Tm, 0, False, True), 0),
            ("x self-attn dense 800x1600", lambda: attn(Tm, 1600, 0, False, False), 0),
            ("x self-attn dense 800x3200", lambda: attn(Tm, 3200, 0, False, False), 0)]

flt = sys.argv[1] if len(sys.argv) > 1 else ""
total = 0.0
for name, mk, count in OPS:
    if flt and flt not in name:
        continue
    fn, flops = mk()
    us = timeit(fn)
    total += us * count
    print(f"{name:52s} {us:8.1f} us x{count}  {flops / us / 1e6:8.1f} TFLOP/s")
    del fn
    torch.cuda.empty_cache()
print(f"sum over one forward: {total:.0f} us")
