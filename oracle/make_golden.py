"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from
/root/reference, CPU, fp32) on seeded weights/inputs, and check the oracle restatement
against it.  Runs only in the build container (the reference cannot travel to the GPU box);
the fixtures it writes are committed.  Test infrastructure, not product.

    python oracle/make_golden.py
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, "/root/reference")
import tts_oracle as O                      # noqa: E402
from model.model import TransformerTTS       # noqa: E402  (the reference)

torch.backends.mha.set_fastpath_enabled(False)   # pin one well-defined torch path
torch.manual_seed(0)
OUT = os.path.join(HERE, "..", "tests", "golden")


def reference_module(sd):
    m = TransformerTTS(**O.CONFIG, device="cpu").eval()
    missing = m.load_state_dict(sd, strict=True)
    return m


def maxdiff(a, b):
    return float((a - b).abs().max())


def fwd_case(name, seed, B, Tp, Tm, keep_align=True):
    sd = O.make_state_dict(seed)
    ref = reference_module(sd)
    ph, mel, pl, ml = O.make_inputs(B, Tp, Tm, seed=seed + 1)
    with torch.no_grad():
        r = ref(ph, mel, pl, ml)
    o = O.forward(sd, ph, mel, pl, ml)
    d = {k: maxdiff(r[k], o[k]) for k in ("pred_melspec", "post_melspec", "pred_stop")}
    d["alignments"] = max(maxdiff(a, b) for a, b in zip(r["alignments"], o["alignments"]))
    print(name, "oracle vs reference max-abs:", d)
    assert max(d.values()) < 2e-4, d
    rec = dict(seed=seed, checksum=O.state_dict_checksum(sd), phoneme=ph.numpy(), melspec=mel.numpy(),
               phoneme_lens=pl.numpy(), melspec_lens=ml.numpy(),
               pred_melspec=r["pred_melspec"].numpy(), post_melspec=r["post_melspec"].numpy(),
               pred_stop=r["pred_stop"].numpy())
    al = torch.stack(r["alignments"]).numpy()
    if keep_align:
        rec["alignments"] = al
    else:   # keep the fixture small: head-averaged maps in fp16 + exact row sums
        rec["alignments_mean_f16"] = al.mean(axis=2).astype(np.float16)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)


def inf_case(name, seed, B, Tp, horizon, min_t):
    sd = O.make_state_dict(seed)
    ph, _, pl, _ = O.make_inputs(B, Tp, 8, seed=seed + 1)
    bias, n_frames, margin = O.plant_stop_bias(sd, ph, pl, horizon, min_t=min_t)
    sd["linear2.linear.bias"] = torch.tensor([bias])
    ref = reference_module(sd)
    r = ref.inference(ph, pl, max_len=horizon + 1)
    o = O.inference(sd, ph, pl, max_len=horizon + 1)
    oc = O.inference_cached(sd, ph, pl, max_len=horizon + 1)
    assert r["pred_melspec"].shape[1] == n_frames == o["pred_melspec"].shape[1] == oc["pred_melspec"].shape[1], \
        (r["pred_melspec"].shape, n_frames)
    for k in ("pred_melspec", "post_melspec", "pred_stop"):
        assert r[k].shape == o[k].shape == oc[k].shape
        print(name, k, "oracle", maxdiff(r[k], o[k]), "oracle_cached", maxdiff(r[k], oc[k]))
        assert maxdiff(r[k], o[k]) < 2e-4 and maxdiff(r[k], oc[k]) < 2e-4
    print(name, "frames", n_frames, "margin", margin, "bias", bias)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), seed=seed, checksum=O.state_dict_checksum(sd),
                        stop_bias=bias, n_frames=n_frames, margin=margin, horizon=horizon,
                        phoneme=ph.numpy(), phoneme_lens=pl.numpy(),
                        pred_melspec=r["pred_melspec"].numpy(), post_melspec=r["post_melspec"].numpy(),
                        pred_stop=r["pred_stop"].numpy())


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    fwd_case("fwd_small", 11, B=3, Tp=12, Tm=40)
    fwd_case("fwd_tile", 12, B=2, Tp=37, Tm=150, keep_align=False)     # crosses a 128-row tile
    fwd_case("fwd_b1", 13, B=1, Tp=9, Tm=17)
    inf_case("inf_b1", 44, B=1, Tp=14, horizon=48, min_t=10)
    inf_case("inf_b3", 38, B=3, Tp=11, horizon=48, min_t=10)         # ragged: encoder unmasked (F8)
    print("golden fixtures written to", os.path.normpath(OUT))
