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
STOP_ATOL_BIG = 0.15     # B >= 64 ragged batches: see test_inference_planted_stop_auto_plan_large_batches


def _model(sd):
    from transformertts_b200 import TransformerTTS
    return TransformerTTS.from_reference(sd, device=DEV)


def _cmp(out, ref, tag, stop_atol=None):
    msgs, ok_all = [], True
    for k, atol, rel in (("pred_melspec", PRE_ATOL, PRE_REL_L2), ("post_melspec", POST_ATOL, POST_REL_L2),
                         ("pred_stop", STOP_ATOL if stop_atol is None else stop_atol, None)):
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


# ------------------------------------------------------------------ the plans behind the configs[3] numbers
# decode_plan() picks R=2 / C=8 (two CTAs per SM) at B=64 and R=2 / C=4 at B=128: those launches are what BASELINE.md's
# 64- and 128-row figures time, so they get the same gates as B=32: fully ragged batches (every row its own utterance),
# planted stop found by oracle search (tools: oracle.plant_stop_bias over seeds), identical stop index, trajectory parity.
@pytest.mark.parametrize("B,Tp,horizon,seed,min_t", [(64, 150, 56, 213, 10), (128, 150, 48, 204, 10)])
def test_inference_planted_stop_auto_plan_large_batches(B, Tp, horizon, seed, min_t):
    from transformertts_b200 import _ext
    assert B <= _ext.load().tts_decode_capacity()            # one wave: the stop rule can end the launch
    sd = O.make_state_dict(seed)
    ph, _, pl, _ = O.make_inputs(B, Tp, 8, seed=seed + 1)
    bias, n_frames, margin = O.plant_stop_bias(sd, ph, pl, horizon, min_t=min_t)
    sd["linear2.linear.bias"] = torch.tensor([bias])
    ref = O.inference_cached(sd, ph, pl, max_len=horizon + 1)
    assert ref["pred_melspec"].shape[1] == n_frames
    m = _model(sd)
    out = m.inference(ph, pl, max_len=horizon + 1)            # rows_per_cluster = cluster_size = 0: the automatic plan
    torch.cuda.synchronize()
    print(f"B={B} auto plan: frames got={out['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert out["pred_melspec"].shape[1] == n_frames
    # The max over 4x more (row, frame) samples than the B=32 case sits further out in the tail of the bf16-feedback
    # error: observed on B200 at B=128 stop 0.132 (1 row of 128, frame 11), pred 0.242 (row 80, frame 12); the stop
    # DECISION is protected by the planted margin (>= 0.25 on both sides), which stays above this tolerance.
    _cmp(out, ref, f"AR B{B} auto plan", stop_atol=STOP_ATOL_BIG)
    # the same rows through the parity-tested 32-row plan: the large-batch plan must not be further from the oracle
    # than a few times the small plan's own error on those rows (a plan bug shows as O(1), chaos as O(0.1))
    lo = 64 if B > 64 else 32
    sub = m.inference(ph[lo:lo + 32], pl[lo:lo + 32], max_len=n_frames + 1, stop_threshold=2.0)   # never stops early
    torch.cuda.synchronize()
    e_big = float((out["pred_melspec"][lo:lo + 32].cpu() - ref["pred_melspec"][lo:lo + 32]).abs().max())
    e_sub = float((sub["pred_melspec"].cpu() - ref["pred_melspec"][lo:lo + 32]).abs().max())
    print(f"rows {lo}-{lo + 31}: pred max_abs vs oracle: {B}-row plan {e_big:.4f}, 32-row plan {e_sub:.4f}")
    assert e_big <= PRE_ATOL and e_sub <= PRE_ATOL


def test_inference_rows_per_cluster_4_runs_and_matches():
    """decode_kernel<4> is selectable through the ABI: run it once (12 ragged rows = 3 clusters of 4) against the oracle."""
    B, Tp, horizon, seed = 12, 50, 40, 144
    sd = O.make_state_dict(seed, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(B, Tp, 8, seed=seed + 1)
    ref = O.inference_cached(sd, ph, pl, max_len=horizon + 1)
    m = _model(sd)
    out = m.inference(ph, pl, max_len=horizon + 1, rows_per_cluster=4)
    auto = m.inference(ph, pl, max_len=horizon + 1)
    torch.cuda.synchronize()
    assert out["pred_melspec"].shape == (B, horizon, 80)
    _cmp(out, ref, "AR R=4")
    _cmp(auto, ref, "AR auto (B=12)")


def test_inference_1000_frames_teacher_forced_b128():
    """The 128-row plan over the full horizon, one step ahead on the oracle's trajectory (the B=32 ragged batch tiled 4x:
    the CPU oracle needs minutes for 128 distinct 1000-frame rows), stop planted at frame 859 in every row."""
    B, rep = 128, 4
    sd, ph, pl, ref = _long_case(32)
    ph, pl = ph.repeat(rep, 1), pl.repeat(rep)
    ref = {k: v.repeat(rep, 1, 1) for k, v in ref.items()}
    out = _model(sd).inference(ph, pl, max_len=1001, forced_melspec=ref["pred_melspec"])
    torch.cuda.synchronize()
    assert out["pred_melspec"].shape == (B, 1000, 80)
    e = float((out["pred_melspec"].float().cpu() - ref["pred_melspec"]).abs().max())
    es = float((out["pred_stop"].float().cpu() - ref["pred_stop"]).abs().max())
    print(f"AR teacher-forced B128: pred max_abs={e:.4f} stop max_abs={es:.4f}")
    assert e <= FORCED_PRE_ATOL and es <= FORCED_STOP_ATOL, (e, es)
    for r in range(1, rep):      # the copies of a row live in different clusters and must agree bit for bit
        assert torch.equal(out["pred_melspec"][:32], out["pred_melspec"][32 * r:32 * (r + 1)])
    sd1, ph1, pl1, ref1 = _long_case(1)
    bias, n_frames, margin = O.plant_from_logits(ref1["pred_stop"][..., 0] + 1.0e4, min_t=LONG_MIN_T, delta=0.25)
    sd2 = dict(sd1)
    sd2["linear2.linear.bias"] = torch.tensor([bias])
    out2 = _model(sd2).inference(ph1.repeat(B, 1), pl1.repeat(B), max_len=1001,
                                 forced_melspec=ref1["pred_melspec"].repeat(B, 1, 1))
    torch.cuda.synchronize()
    print(f"AR teacher-forced planted B128: frames got={out2['pred_melspec'].shape[1]} ref={n_frames} margin={margin:.3f}")
    assert out2["pred_melspec"].shape[1] == n_frames
    assert float((out2["pred_melspec"][:1].cpu() - ref1["pred_melspec"][:, :n_frames]).abs().max()) <= FORCED_PRE_ATOL


# ------------------------------------------------------------------ sharding.sharded_inference on the real module
@pytest.mark.parametrize("micro_batch", [32, 64])
def test_sharded_inference_equals_direct_calls_and_the_oracle(micro_batch):
    """96 ragged utterances dealt into micro-batches: every micro-batch result equals a direct inference() call on the
    same rows bit for bit, and the oracle (run per micro-batch: the all-rows stop rule couples the rows of ONE call,
    model/model.py:382-384) gives the same frame count and trajectory."""
    from transformertts_b200.sharding import micro_batches, sharded_inference
    N, Tp, horizon, seed = 96, 60, 24, 150
    sd = O.make_state_dict(seed, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(N, Tp, 8, seed=seed + 1)
    m = _model(sd)
    parts = sharded_inference(m, ph, pl, micro_batch=micro_batch, max_len=horizon + 1, rank=0, world_size=1)
    torch.cuda.synchronize()
    assert [(s, e) for s, e, _ in parts] == micro_batches(0, N, micro_batch)
    for s, e, res in parts:
        tp = int(pl[s:e].max())
        direct = m.inference(ph[s:e, :tp], pl[s:e], max_len=horizon + 1)
        ref = O.inference_cached(sd, ph[s:e, :tp], pl[s:e], max_len=horizon + 1)
        torch.cuda.synchronize()
        for k in ("pred_melspec", "post_melspec", "pred_stop"):
            assert torch.equal(res[k], direct[k]), (s, e, k)
        assert res["pred_melspec"].shape[1] == ref["pred_melspec"].shape[1]
        _cmp(res, ref, f"sharded rows {s}-{e}")
    # two ranks, no process group needed: rank r owns a contiguous block, together they cover every utterance once
    a = sharded_inference(m, ph, pl, micro_batch=micro_batch, max_len=horizon + 1, rank=0, world_size=2)
    b = sharded_inference(m, ph, pl, micro_batch=micro_batch, max_len=horizon + 1, rank=1, world_size=2)
    cover = sorted((s, e) for s, e, _ in a + b)
    assert cover[0][0] == 0 and cover[-1][1] == N and all(x[1] == y[0] for x, y in zip(cover, cover[1:]))
