"""Host-side profile of harness.LightningModule.validation_step under graph replay (per-step .item() makes the host
path part of the step).  Usage: python tools/profile_validation_step.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, ROOT + "/oracle")
import tts_oracle as O
from transformertts_b200.data import PinnedFeeder
from transformertts_b200.harness import LightningModule

cfg = {"model": {}, "loss": {"stop_weight": 8.0}, "training": {"num_epochs": 300, "teacher_forcing_mode": "linear",
                                                               "log_interval": 10 ** 9}}
lm = LightningModule(cfg, device="cuda:0")
lm.model.use_cuda_graph = "--no-graph" not in sys.argv
ph, mel, pl, ml = O.make_inputs(64, 150, 800, seed=1)
batch = {"phoneme": ph.pin_memory(), "melspec": mel.pin_memory(), "phoneme_lens": pl.pin_memory(), "melspec_lens": ml.pin_memory()}


def run(n):
    for i, b in enumerate(PinnedFeeder((batch for _ in range(n)), "cuda:0")):
        lm.validation_step(b, i + 1)
    torch.cuda.synchronize()


run(5)
t0 = time.perf_counter(); run(200); dt = (time.perf_counter() - t0) / 200
print(f"validation_step: {dt * 1e3:.3f} ms per step")
pr = cProfile.Profile(); pr.enable(); run(200); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
