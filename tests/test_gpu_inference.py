"""GPU parity of the KV-cached AR decode (one persistent kernel) vs the reference's inference():
identical stop-frame index on planted-stop cases, trajectory drift within tolerance, and the
quirks SURVEY.md lists (F8 unmasked encoder, F9 all-rows stop rule / (B,T,1) stop shape / max_len-1 frames)."""
import os

import numpy as np
import pytest
import torch

import tts_oracle as O
from parity_common import DEV, report

pytestmark = pytest.mark.gpu

# Stated AR tolerance.  The fed-back frame re-enters the decoder, so bf16 rounding (K/V cache, PostNet
# activations) compounds along the trajectory; measured on B200 (round 1): pred max-abs <= 0.16,
# post <= 0.25 (tail: < 0.3 % of elements above 0.15), relative L2 <= 1.2e-2 / 1.9e-2.  The hard gate is the
# identical stop-frame index, asserted separately in every test below.
PRE_ATOL, POST_ATOL, STOP_ATOL, PRE_REL_L2, POST_REL_L2 = 0.25, 0.35, 0.10, 2e-2, 3e-2


def _model(sd):
    from transformertts_b200 import TransformerTTS
    return TransformerTTS.from_reference(sd, device=DEV)


def _cmp(out, ref, tag):
    msgs, ok_all = [], True
    for k, atol, rel in (("pred_melspec", PRE_ATOL, PRE_REL_L2), ("post_melspec", POST_ATOL, POST_REL_L2),
                         ("pred_stop", STOP_ATOL, None)):
        ok, msg = report(f"{tag} {k}", out[k], torch.as_tensor(ref[k]), atol)
        msgs.append(msg); ok_all &= ok
        if rel is not None:
            g, r = out[k].float().cpu(), torch.as_tensor(ref[k]).float()
            ok_all &= float((g - r).norm() / r.norm()) <= rel
    print("\n".join(msgs))
    assert ok_all, "\n".join(msgs)


@pytest.mark.parametrize("name", ["inf_b1", "inf_b3"])
@pytest.mark.parametrize("cluster", [16, 8])
def test_inference_vs_reference_golden(golden_dir, name, cluster):
    g = dict(np.load(os.path.join(golden_dir, name + ".npz")))
    sd = O.make_state_dict(int(g["seed"]))
    sd["linear2.linear.bias"] = torch.tensor([float(g["stop_bias"])])
    m = _model(sd)
    out = m.inference(torch.from_numpy(g["phoneme"]), torch.from_numpy(g["phoneme_lens"]),
                      max_len=int(g["horizon"]) + 1, cluster_size=cluster)
    torch.cuda.synchronize()
    print(f"{name}: frames got={out['pred_melspec'].shape[1]} ref={int(g['n_frames'])} margin={float(g['margin']):.3f}")
    assert out["pred_melspec"].shape[1] == int(g["n_frames"])          # identical stop-frame index
    assert out["pred_stop"].shape == g["pred_stop"].shape              # (B,T,1)
    _cmp(out, g, f"{name}/c{cluster}")


# seeds found by searching for a trajectory whose all-rows stop logit makes a late new maximum with margin
# >= 0.25 on both sides (oracle.plant_stop_bias); B=32 is BASELINE.json config 3's batch
@pytest.mark.parametrize("B,Tp,horizon,rows,seed,min_t", [(1, 150, 120, 0, 100, 20), (5, 60, 90, 0, 126, 15),
                                                          (32, 150, 64, 0, 143, 12), (6, 40, 50, 2, 144, 10),
                                                          (9, 30, 40, 8, 144, 10)])
def test_inference_planted_stop_vs_oracle(B, Tp, horizon, rows, seed, min_t):
    sd = O.make_state_dict(seed)
    ph, _, pl, _ = O.make_inputs(B, Tp, 8, seed=seed + 1)
    bias, n_frames, margin = O.plant_stop_bias(sd, ph, pl, horizon, min_t=min_t)
    sd["linear2.linear.bias"] = torch.tensor([bias])
    ref = O.inference_cached(sd, ph, pl, max_len=horizon + 1)
    assert ref["pred_melspec"].shape[1] == n_frames
    m = _model(sd)
    out = m.inference(ph, pl, max_len=horizon + 1, rows_per_cluster=rows)
    torch.cuda.synchronize()
    print(f"B={B}: frames got={out['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert out["pred_melspec"].shape[1] == n_frames
    _cmp(out, ref, f"AR B{B}")


def test_inference_never_stops_hits_max_len():
    sd = O.make_state_dict(70, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(2, 20, 8, seed=71)
    m = _model(sd)
    out = m.inference(ph, pl, max_len=25)
    assert out["pred_melspec"].shape == (2, 24, 80) and out["pred_stop"].shape == (2, 24, 1)   # max_len-1 frames
    ref = O.inference_cached(sd, ph, pl, max_len=25)
    _cmp(out, ref, "never-stop")


def test_inference_immediate_stop():
    sd = O.make_state_dict(72, stop_bias=1e4)
    ph, _, pl, _ = O.make_inputs(3, 20, 8, seed=73)
    out = _model(sd).inference(ph, pl, max_len=50)
    assert out["pred_melspec"].shape == (3, 1, 80)


def test_inference_matches_forward_on_unpadded_batch():
    """AR decode == teacher-forced forward over its own output for unpadded batches;
    (iv): it must differ on padded batches because inference() leaves the encoder unmasked (F8)."""
    sd = O.make_state_dict(74, stop_bias=-1e4)
    m = _model(sd)
    ph, _, pl, _ = O.make_inputs(2, 24, 8, seed=75, ragged=False)
    inf = m.inference(ph, pl, max_len=41)
    T = inf["pred_melspec"].shape[1]
    fwd = m(ph, inf["pred_melspec"], pl, torch.full((2,), T))
    ok, msg = report("AR vs forward(unpadded)", fwd["pred_melspec"], inf["pred_melspec"], 0.12)
    print(msg)
    assert ok, msg
    ph2, _, pl2, _ = O.make_inputs(3, 24, 8, seed=76, ragged=True)
    assert int(pl2.min()) < 24
    inf2 = m.inference(ph2, pl2, max_len=41)
    fwd2 = m(ph2, inf2["pred_melspec"], pl2, torch.full((3,), inf2["pred_melspec"].shape[1]))
    assert float((fwd2["pred_melspec"] - inf2["pred_melspec"]).abs().max()) > 0.2


# ------------------------------------------------------------------ the benchmarked horizon: 1000 frames (configs[2])
# One never-stopping oracle trajectory per batch size serves both tests below: the stop logit is not fed back, so the
# planted-stop run is the same trajectory cut at the planted step (PostNet recomputed over the cut prefix).
_LONG = {}
LONG_SEED, LONG_MIN_T = 221, 800     # found by oracle search: the stop logit makes a new maximum at frame 859, margin 0.34
# Observed drift at 1000 frames is printed per 100-frame bucket; gates are the AR tolerances above.


def _long_case(B, replicate=False):
    """replicate: all B rows are the B=1 utterance (a late all-rows crossing with margin does not occur by chance for 32
    independent rows: the minimum over rows of 32 random logit tracks has no late record); the rows still go through the
    batch-32 decode plan and its cross-cluster stop counting.  The drift test uses a real ragged batch."""
    key = (B, replicate)
    if key not in _LONG:
        sd = O.make_state_dict(LONG_SEED, stop_bias=-1e4)
        if replicate:
            ph, _, pl, _ = O.make_inputs(1, 150, 8, seed=LONG_SEED + 1)
            ph, pl = ph.repeat(B, 1), pl.repeat(B)
            one = _long_case(1)[3]
            ref = {k: v.repeat(B, 1, 1) for k, v in one.items()}
        else:
            ph, _, pl, _ = O.make_inputs(B, 150, 8, seed=LONG_SEED + 1)
            ref = O.inference_cached(sd, ph, pl, max_len=1001)
        assert ref["pred_melspec"].shape[1] == 1000
        _LONG[key] = (sd, ph, pl, ref)
    return _LONG[key]


@pytest.mark.parametrize("B", [1, 32])
def test_inference_1000_frames_never_stop_drift(B):
    sd, ph, pl, ref = _long_case(B)
    out = _model(sd).inference(ph, pl, max_len=1001)
    torch.cuda.synchronize()
    assert out["pred_melspec"].shape == (B, 1000, 80)
    g, r = out["pred_melspec"].float().cpu(), ref["pred_melspec"]
    gp, rp = out["post_melspec"].float().cpu(), ref["post_melspec"]
    for s in range(0, 1000, 100):
        e = (g[:, s:s + 100] - r[:, s:s + 100])
        ep = (gp[:, s:s + 100] - rp[:, s:s + 100])
        print(f"AR B{B} frames {s:4d}-{s + 99:4d}: pred max_abs={float(e.abs().max()):.4f} "
              f"rel_l2={float(e.norm() / r[:, s:s + 100].norm()):.4f} | post max_abs={float(ep.abs().max()):.4f} "
              f"rel_l2={float(ep.norm() / rp[:, s:s + 100].norm()):.4f} | stop max_abs="
              f"{float((out['pred_stop'].cpu()[:, s:s + 100] - ref['pred_stop'][:, s:s + 100]).abs().max()):.4f}")
    _cmp(out, ref, f"AR-1000 never-stop B{B}")


@pytest.mark.parametrize("B", [1, 32])
def test_inference_1000_frame_horizon_planted_stop_index(B):
    """Identical stop-frame index (model/model.py:382-384) late in the 1000-frame horizon, margin >= 0.25 on both sides."""
    sd, ph, pl, ref = _long_case(B, replicate=B > 1)
    z = ref["pred_stop"][..., 0] + 1.0e4
    bias, n_frames, margin = O.plant_from_logits(z, min_t=LONG_MIN_T, delta=0.25)
    sd2 = dict(sd)
    sd2["linear2.linear.bias"] = torch.tensor([bias])
    out = _model(sd2).inference(ph, pl, max_len=1001)
    torch.cuda.synchronize()
    print(f"AR-1000 planted B{B}: frames got={out['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert n_frames >= LONG_MIN_T and margin >= 0.25
    assert out["pred_melspec"].shape[1] == n_frames                       # identical stop-frame index
    pred = ref["pred_melspec"][:, :n_frames]
    cut = {"pred_melspec": pred, "post_melspec": O.postnet(sd, pred) + pred,
           "pred_stop": ref["pred_stop"][:, :n_frames] + 1.0e4 + bias}
    _cmp(out, cut, f"AR-1000 planted B{B}")
