"""Two teacher-forced forwards at BASELINE configs[1] (B=64, Tp=150, Tm=800) for ncu captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import make_batch
from transformertts_b200 import TransformerTTS

B = int(os.environ.get("PB", 64))
m = TransformerTTS(device="cuda:0").eval()
ph, mel, pl, ml = (t.cuda() for t in make_batch(B, 150, 800, 1))
for _ in range(2):
    out = m(ph, mel, pl, ml)
torch.cuda.synchronize()
print("ok", float(out["post_melspec"].abs().mean()))
