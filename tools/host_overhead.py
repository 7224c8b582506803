"""Host-side cost of one forward call vs its device time (cfg 2 and small shapes): if enqueueing ~50 launches takes
longer than the GPU needs to run them, the step is host bound and a CUDA graph pays.  Usage: python tools/host_overhead.py"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transformertts_b200 import TransformerTTS
from bench import make_batch

dev = torch.device("cuda:0")
m = TransformerTTS(device=dev).eval()
for (B, Tp, Tm) in ((64, 150, 800), (16, 150, 800), (1, 400, 2000), (1, 150, 800)):
    x = [t.to(dev) for t in make_batch(B, Tp, Tm, seed=1)]
    for _ in range(5):
        m(*x)
    torch.cuda.synchronize()
    n = 50
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(n):
        m(*x)
    t_enq = time.perf_counter() - t0
    b.record()
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print(f"B={B} Tp={Tp} Tm={Tm}: host enqueue {1e3 * t_enq / n:.3f} ms/forward, device {a.elapsed_time(b) / n:.3f} ms/forward, "
          f"wall {1e3 * t_all / n:.3f} ms/forward")
