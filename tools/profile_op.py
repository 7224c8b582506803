"""A few launches of one operator shape of tools/time_ops.py for ncu (no timing).  Usage: profile_op.py <substring>"""
import os, sys
os.environ["TTS_TIME_OPS_WARM"] = "1"
flt = sys.argv[1]
sys.argv = [sys.argv[0], "\0nothing"]          # import time_ops without running its table
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import time_ops as T
import torch
for name, mk, _ in T.OPS:
    if flt in name:
        fn, _ = mk()
        for _ in range(4):
            fn()
        torch.cuda.synchronize()
        print("ran", name)
