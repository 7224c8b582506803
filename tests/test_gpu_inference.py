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


def test_inference_multi_wave_launch_keeps_the_stop_index():
    """A batch larger than one launch holds co-resident (here forced: 40 one-row clusters of 16 CTAs = 640 CTAs) runs in
    waves; the first wave cannot see the later clusters report, so it decodes to max_frames, but the finalize pass over
    the per-step counters must still return the reference's stop index."""
    from transformertts_b200 import _ext
    sd = O.make_state_dict(100, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(1, 150, 8, seed=101)
    ref = O.inference_cached(sd, ph, pl, max_len=61)
    bias, n_frames, margin = O.plant_from_logits(ref["pred_stop"][..., 0] + 1.0e4, min_t=20, delta=0.25)
    sd2 = dict(sd)
    sd2["linear2.linear.bias"] = torch.tensor([bias])
    B = 40
    assert B * 16 > _ext.load().tts_decode_max_clusters(1, 16) * 16
    out = _model(sd2).inference(ph.repeat(B, 1), pl.repeat(B), max_len=61, rows_per_cluster=1, cluster_size=16)
    torch.cuda.synchronize()
    print(f"multi-wave: frames got={out['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert out["pred_melspec"].shape == (B, n_frames, 80)
    assert torch.equal(out["pred_melspec"][0], out["pred_melspec"][B - 1])
    assert float((out["pred_melspec"][0].cpu() - ref["pred_melspec"][0, :n_frames]).abs().max()) <= PRE_ATOL


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
# At random init the AR map is chaotic over this horizon: the REFERENCE ITSELF, with its encoder output perturbed by
# 1e-6, is 2e-2 away from its own trajectory at frame 1000, and 1.6 away for a 1e-4 perturbation (measured with the
# oracle; table in BASELINE.md).  A bf16 K/V cache (2^-9 relative rounding) therefore cannot stay on the fp32
# trajectory for 1000 free-running frames, and no implementation that rounds anywhere can.  So the full horizon is
# pinned in two ways:
#   * free-running: the drift is printed per 100-frame bucket for B=1 / B=32 and gated where it is still rounding-level
#     for every row (the rows of a batch lose the reference trajectory at different times: the B=1 utterance stays within
#     0.03 for 400 frames, some rows of the ragged B=32 batch leave after ~80); a free-running planted stop at frame 356
#     (B=1) must give the identical stop index and trajectory;
#   * teacher-forced decode (tts_decode's `forced_mel`): every one of the 1000 steps runs through the same persistent
#     kernel and K/V cache but consumes the ORACLE's previous frame, so errors cannot compound through the feedback:
#     all 1000 frames and stop logits are compared one step ahead, and a stop planted at frame 859 of the horizon
#     (margin 0.34) must be found at the identical index.
_LONG = {}
LONG_SEED, LONG_MIN_T = 221, 800     # found by oracle search: the stop logit makes a new maximum at frame 859, margin 0.34
FREE_GATE_FRAMES = {1: 300, 32: 100}  # free-running frames gated at the AR tolerances above
# one-step-ahead over 1000 frames (bf16 K/V cache + bf16 cross K/V): observed on B200 pred 0.0093 / stop 0.0059 / post 0.075
# (ragged B=32 batch; B=1: 0.0057 / 0.0039)
FORCED_PRE_ATOL, FORCED_STOP_ATOL, FORCED_POST_ATOL = 0.014, 0.009, 0.11


def _long_case(B):
    if B not in _LONG:
        sd = O.make_state_dict(LONG_SEED, stop_bias=-1e4)
        ph, _, pl, _ = O.make_inputs(B, 150, 8, seed=LONG_SEED + 1)
        ref = O.inference_cached(sd, ph, pl, max_len=1001)
        assert ref["pred_melspec"].shape[1] == 1000
        _LONG[B] = (sd, ph, pl, ref)
    return _LONG[B]


@pytest.mark.parametrize("B", [1, 32])
def test_inference_1000_frames_free_running_drift(B):
    sd, ph, pl, ref = _long_case(B)
    out = _model(sd).inference(ph, pl, max_len=1001)
    torch.cuda.synchronize()
    assert out["pred_melspec"].shape == (B, 1000, 80) and out["pred_stop"].shape == (B, 1000, 1)
    g, r = out["pred_melspec"].float().cpu(), ref["pred_melspec"]
    gp, rp = out["post_melspec"].float().cpu(), ref["post_melspec"]
    for s in range(0, 1000, 100):
        e = (g[:, s:s + 100] - r[:, s:s + 100])
        ep = (gp[:, s:s + 100] - rp[:, s:s + 100])
        print(f"AR free-running B{B} frames {s:4d}-{s + 99:4d}: pred max_abs={float(e.abs().max()):.4f} "
              f"rel_l2={float(e.norm() / r[:, s:s + 100].norm()):.4f} | post max_abs={float(ep.abs().max()):.4f} "
              f"rel_l2={float(ep.norm() / rp[:, s:s + 100].norm()):.4f} | stop max_abs="
              f"{float((out['pred_stop'].cpu()[:, s:s + 100] - ref['pred_stop'][:, s:s + 100]).abs().max()):.4f}")
    assert torch.isfinite(out["post_melspec"]).all()
    n = FREE_GATE_FRAMES[B]
    # PostNet has a 5x5-tap receptive field of +-10 frames: compare post over the gated prefix minus that halo
    head = {"pred_melspec": out["pred_melspec"][:, :n], "post_melspec": out["post_melspec"][:, :n - 10],
            "pred_stop": out["pred_stop"][:, :n]}
    href = {"pred_melspec": ref["pred_melspec"][:, :n], "post_melspec": ref["post_melspec"][:, :n - 10],
            "pred_stop": ref["pred_stop"][:, :n]}
    if B == 1:
        _cmp(head, href, f"AR-1000 free-running B{B}, first {n} frames")
    else:
        # a ragged batch always holds a few rows that leave the reference trajectory within tens of frames (row 28 of this
        # one is 0.40 off at frame 29): gate the batch by relative L2 over the prefix, the rows by their median max-abs
        for k, rel in (("pred_melspec", PRE_REL_L2), ("post_melspec", POST_REL_L2)):
            gg, rr = head[k].float().cpu(), href[k]
            r2 = float((gg - rr).norm() / rr.norm())
            med = float((gg - rr).abs().flatten(1).max(dim=1).values.median())
            print(f"AR-1000 free-running B{B}, first {n} frames {k}: rel_l2={r2:.4f} median-row max_abs={med:.4f}")
            assert r2 <= rel and med <= (PRE_ATOL if k == "pred_melspec" else POST_ATOL), (k, r2, med)


@pytest.mark.parametrize("B,seed,min_t", [(1, 317, 250)])
def test_inference_free_running_planted_stop_late(B, seed, min_t):
    """Identical stop-frame index (model/model.py:382-384) for a free-running decode as late as the chaos allows."""
    sd = O.make_state_dict(seed, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(B, 150, 8, seed=seed + 1)
    ref = O.inference_cached(sd, ph, pl, max_len=401)
    bias, n_frames, margin = O.plant_from_logits(ref["pred_stop"][..., 0] + 1.0e4, min_t=min_t, delta=0.25)
    sd2 = dict(sd)
    sd2["linear2.linear.bias"] = torch.tensor([bias])
    out = _model(sd2).inference(ph, pl, max_len=401)
    torch.cuda.synchronize()
    print(f"AR free-running planted B{B}: frames got={out['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert n_frames >= min_t and out["pred_melspec"].shape[1] == n_frames
    pred = ref["pred_melspec"][:, :n_frames]
    cut = {"pred_melspec": pred, "post_melspec": O.postnet(sd, pred) + pred,
           "pred_stop": ref["pred_stop"][:, :n_frames] + 1.0e4 + bias}
    _cmp(out, cut, f"AR free-running planted B{B}")


@pytest.mark.parametrize("B", [1, 32])
def test_inference_1000_frames_teacher_forced_parity_and_stop_index(B):
    """All 1000 steps through the decode kernel and its K/V cache on the oracle's trajectory: one-step-ahead parity of
    every frame and stop logit, and the identical stop index for a stop planted at frame 859 (B=32: the B=1 utterance in
    every row -- 32 independent rows have no late all-rows crossing -- through the batch-32 plan and its cross-cluster
    stop counting)."""
    sd, ph, pl, ref = _long_case(B)          # B=32: the ragged batch (encoder unmasked over the padding)
    m = _model(sd)
    out = m.inference(ph, pl, max_len=1001, forced_melspec=ref["pred_melspec"])
    torch.cuda.synchronize()
    assert out["pred_melspec"].shape == (B, 1000, 80)
    e = (out["pred_melspec"].float().cpu() - ref["pred_melspec"]).abs()
    es = (out["pred_stop"].float().cpu() - ref["pred_stop"]).abs()
    for s in range(0, 1000, 100):
        print(f"AR teacher-forced B{B} frames {s:4d}-{s + 99:4d}: pred max_abs={float(e[:, s:s + 100].max()):.4f} "
              f"stop max_abs={float(es[:, s:s + 100].max()):.4f}")
    assert float(e.max()) <= FORCED_PRE_ATOL and float(es.max()) <= FORCED_STOP_ATOL, (float(e.max()), float(es.max()))
    ok, msg = report(f"AR teacher-forced B{B} post", out["post_melspec"], ref["post_melspec"], FORCED_POST_ATOL)
    print(msg)
    assert ok, msg
    # planted stop late in the horizon
    sd, ph, pl, ref = _long_case(1)
    if B > 1:
        ph, pl = ph.repeat(B, 1), pl.repeat(B)
        ref = {k: v.repeat(B, 1, 1) for k, v in ref.items()}
    bias, n_frames, margin = O.plant_from_logits(ref["pred_stop"][..., 0] + 1.0e4, min_t=LONG_MIN_T, delta=0.25)
    sd2 = dict(sd)
    sd2["linear2.linear.bias"] = torch.tensor([bias])
    out2 = _model(sd2).inference(ph, pl, max_len=1001, forced_melspec=ref["pred_melspec"])
    torch.cuda.synchronize()
    print(f"AR teacher-forced planted B{B}: frames got={out2['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert n_frames >= LONG_MIN_T and margin >= 0.25
    assert out2["pred_melspec"].shape[1] == n_frames                      # identical stop-frame index
    assert float((out2["pred_melspec"].cpu() - ref["pred_melspec"][:, :n_frames]).abs().max()) <= FORCED_PRE_ATOL
    if B > 1:       # identical rows stay bit-identical across the clusters that own them
        assert torch.equal(out2["pred_melspec"][0], out2["pred_melspec"][B - 1])
