"""Soak run of the CUDA path: many teacher-forced forwards (and a few AR decodes) over random shapes and lengths,
each forward executed twice and compared bit for bit.  The kernels hand work between roles through mbarrier
protocols; a protocol race shows up either as a trapped kernel (bounded waits, ptx.cuh) or as two runs of the same
input that differ.  Usage: python tools/soak.py [iterations] [seed]      (run it under `timeout`)"""
import os
import random
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import make_batch
from transformertts_b200 import TransformerTTS

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rng = random.Random(seed)
m = TransformerTTS(device="cuda:0").eval()
t0 = time.time()
# SOAK_NOISE=1: unrelated kernels of random size on a side stream while the forwards run, to perturb CTA placement
# and timing (protocol races are timing dependent)
noise = os.environ.get("SOAK_NOISE")
ns = torch.cuda.Stream() if noise else None
nbuf = torch.zeros(64 << 20, dtype=torch.float32, device="cuda:0") if noise else None
shapes = [(64, 150, 800), (16, 150, 800), (1, 33, 130), (2, 400, 1000), (3, 64, 257), (64, 150, 800)]
for i in range(iters):
    if i < len(shapes):
        B, Tp, Tm = shapes[i]
    else:
        B, Tp, Tm = rng.choice([1, 2, 3, 5, 8, 16, 33, 64]), rng.randint(8, 260), rng.randint(16, 1100)
    ph, mel, pl, ml = (t.cuda() for t in make_batch(B, Tp, Tm, seed * 100003 + i))
    want_align = rng.random() < 0.7
    if noise:
        with torch.cuda.stream(ns):
            for _ in range(rng.randint(1, 6)):
                n = rng.randint(1 << 10, 64 << 20)
                nbuf[:n].mul_(1.0001).add_(0.5)
    kw = {} if want_align else {"return_alignments": False}
    try:
        a = m(ph, mel, pl, ml, **kw)
        b = m(ph, mel, pl, ml, **kw)
    except TypeError:
        a = m(ph, mel, pl, ml)
        b = m(ph, mel, pl, ml)
    torch.cuda.synchronize()
    for k in ("pred_melspec", "post_melspec", "pred_stop"):
        assert torch.isfinite(a[k]).all(), f"iteration {i} {k}: non-finite output at shape {(B, Tp, Tm)}"
        assert torch.equal(a[k], b[k]), f"iteration {i} {k}: two runs differ at shape {(B, Tp, Tm)}"
    if a.get("alignments") is not None:
        for x, y in zip(a["alignments"], b["alignments"]):
            assert torch.equal(x, y), f"iteration {i}: alignments of two runs differ at shape {(B, Tp, Tm)}"
    if i % 25 == 24:
        Bi = rng.choice([1, 4, 32])
        iph, _, ipl, _ = make_batch(Bi, rng.randint(20, 150), 8, seed + i)
        ml_ = rng.randint(40, 200)
        r1 = m.inference(iph.cuda(), ipl.cuda(), max_len=ml_)
        r2 = m.inference(iph.cuda(), ipl.cuda(), max_len=ml_)
        torch.cuda.synchronize()
        assert torch.isfinite(r1["pred_melspec"]).all()
        for k in r1:
            assert torch.equal(r1[k], r2[k]), f"iteration {i}: AR decode {k} of two runs differs (B={Bi})"
    if i % 20 == 19:
        print(f"{i + 1} iterations, {time.time() - t0:.0f} s", flush=True)
print(f"soak ok: {iters} iterations in {time.time() - t0:.0f} s")
