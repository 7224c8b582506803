"""CPU tests: the oracle restatement against the golden vectors produced by the real
reference (oracle/make_golden.py), plus the properties SURVEY.md section 4 lists."""
import os

import numpy as np
import pytest
import torch

import tts_oracle as O

FWD = ["fwd_small", "fwd_tile", "fwd_b1"]
INF = ["inf_b1", "inf_b3"]


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


@pytest.mark.parametrize("name", FWD)
def test_forward_matches_reference_golden(golden_dir, name):
    g = _load(golden_dir, name)
    sd = O.make_state_dict(int(g["seed"]))
    assert abs(O.state_dict_checksum(sd) - float(g["checksum"])) < 1e-6 * float(g["checksum"])
    out = O.forward(sd, torch.from_numpy(g["phoneme"]), torch.from_numpy(g["melspec"]),
                    torch.from_numpy(g["phoneme_lens"]), torch.from_numpy(g["melspec_lens"]))
    for k in ("pred_melspec", "post_melspec", "pred_stop"):
        assert out[k].shape == g[k].shape
        np.testing.assert_allclose(out[k].numpy(), g[k], atol=2e-4, rtol=0)
    al = torch.stack(out["alignments"]).numpy()
    if "alignments" in g:
        np.testing.assert_allclose(al, g["alignments"], atol=1e-5, rtol=0)
    else:
        np.testing.assert_allclose(al.mean(axis=2), g["alignments_mean_f16"].astype(np.float32), atol=1e-3)


@pytest.mark.parametrize("name", INF)
def test_inference_matches_reference_golden(golden_dir, name):
    g = _load(golden_dir, name)
    sd = O.make_state_dict(int(g["seed"]))
    sd["linear2.linear.bias"] = torch.tensor([float(g["stop_bias"])])
    ph, pl = torch.from_numpy(g["phoneme"]), torch.from_numpy(g["phoneme_lens"])
    for fn in (O.inference, O.inference_cached):
        out = fn(sd, ph, pl, max_len=int(g["horizon"]) + 1)
        assert out["pred_melspec"].shape[1] == int(g["n_frames"])
        assert out["pred_stop"].shape == g["pred_stop"].shape        # (B,T,1)  SURVEY F9
        for k in ("pred_melspec", "post_melspec", "pred_stop"):
            np.testing.assert_allclose(out[k].numpy(), g[k], atol=2e-4, rtol=0)


def test_alignment_rows_sum_to_one_and_zero_on_padding():
    sd = O.make_state_dict(5)
    ph, mel, pl, ml = O.make_inputs(3, 13, 21, seed=6)
    out = O.forward(sd, ph, mel, pl, ml)
    for A in out["alignments"]:
        assert A.shape == (3, 4, 21, 13)
        np.testing.assert_allclose(A.sum(-1).numpy(), 1.0, atol=1e-5)
        for b in range(3):
            assert float(A[b, :, :, int(pl[b]):].abs().max()) == 0.0 if int(pl[b]) < 13 else True


def test_max_len_caps_frames():
    sd = O.make_state_dict(7, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(2, 6, 8, seed=8)
    out = O.inference_cached(sd, ph, pl, max_len=5)
    assert out["pred_melspec"].shape == (2, 4, 80)      # at most max_len-1 frames (F9)


def test_inference_encoder_is_unmasked():
    """inference() runs the encoder without the key-padding mask; forward() masks.
    AR output fed back through forward() therefore agrees only for unpadded batches."""
    sd = O.make_state_dict(9, stop_bias=-1e4)
    ph, _, pl, _ = O.make_inputs(2, 9, 8, seed=10, ragged=False)
    inf = O.inference_cached(sd, ph, pl, max_len=13)
    T = inf["pred_melspec"].shape[1]
    fwd = O.forward(sd, ph, inf["pred_melspec"], pl, torch.full((2,), T))
    np.testing.assert_allclose(fwd["pred_melspec"].numpy(), inf["pred_melspec"].numpy(), atol=1e-4)
    ph2, _, pl2, _ = O.make_inputs(3, 9, 8, seed=11, ragged=True)
    assert int(pl2.min()) < 9
    inf2 = O.inference_cached(sd, ph2, pl2, max_len=13)
    fwd2 = O.forward(sd, ph2, inf2["pred_melspec"], pl2, torch.full((3,), inf2["pred_melspec"].shape[1]))
    assert float((fwd2["pred_melspec"] - inf2["pred_melspec"]).abs().max()) > 1e-2
