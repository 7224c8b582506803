"""One AR decode (B utterances, max_len frames) for ncu / timing. Usage: profile_decode.py B max_len [R] [C]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import make_batch
from transformertts_b200 import TransformerTTS
B, L = int(sys.argv[1]), int(sys.argv[2])
R = int(sys.argv[3]) if len(sys.argv) > 3 else 0
C = int(sys.argv[4]) if len(sys.argv) > 4 else 0
torch.manual_seed(0)
m = TransformerTTS(device="cuda:0").eval()
with torch.no_grad():
    m.linear2.linear.bias.fill_(-1e4)
m.refresh_weights()
ph, _, pl, _ = make_batch(B, 150, 8, 3)
ph, pl = ph.cuda(), pl.cuda()
for i in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = m.inference(ph, pl, max_len=L, rows_per_cluster=R, cluster_size=C)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
T = out["pred_melspec"].shape[1]
print(f"B={B} frames={T} R={R} C={C}: {dt*1e3:.2f} ms, {dt*1e6/T:.1f} us/step, {B*T/dt:.0f} frames/s")
