"""Probe: does replaying the forward as a CUDA graph (torch.cuda.CUDAGraph around the ctypes launches) beat stream
launches?  Small shapes are launch/latency bound (50 launches).  Usage: python tools/graph_probe.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import make_batch
from transformertts_b200 import TransformerTTS

dev = torch.device("cuda:0")
m = TransformerTTS(device=dev).eval()
for (B, Tp, Tm) in ((1, 400, 2000), (1, 150, 800), (16, 150, 800), (64, 150, 800)):
    x = [t.to(dev) for t in make_batch(B, Tp, Tm, seed=1)]
    for _ in range(3):
        ref = m(*x)
    torch.cuda.synchronize()

    def timeit(fn, n=30):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn()
        b.record(); b.synchronize()
        return a.elapsed_time(b) / n
    t_stream = timeit(lambda: m(*x))
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream(dev)
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        m(*x)
    torch.cuda.current_stream().wait_stream(s)
    try:
        with torch.cuda.graph(g):
            out = m(*x)
        g.replay(); torch.cuda.synchronize()
        same = all(torch.equal(out[k], ref[k]) for k in ("pred_melspec", "post_melspec", "pred_stop"))
        t_graph = timeit(g.replay)
        print(f"B={B} Tp={Tp} Tm={Tm}: stream {t_stream:.3f} ms, graph {t_graph:.3f} ms, identical={same}")
    except Exception as e:
        print(f"B={B} Tp={Tp} Tm={Tm}: stream {t_stream:.3f} ms, graph capture failed: {type(e).__name__}: {str(e)[:200]}")
