"""GPU parity of the forward-only training / validation pieces through the C ABI:
scheduled-sampling mix and loss kernels vs the reference-pinned oracle (oracle/train_oracle.py) and the reference's own
golden outputs, the pinned feeder / fetcher round trip, bf16 mel input, and the Lightning-free harness running the
reference's validation_step / on_validation_epoch_end / training_step bodies on the B200 module."""
import os

import numpy as np
import pytest
import torch

import train_oracle as TO
import tts_oracle as O
from parity_common import DEV

pytestmark = pytest.mark.gpu

CONFIG = {"model": dict(O.CONFIG), "loss": {"stop_weight": 8.0},
          "training": {"num_epochs": 300, "teacher_forcing_mode": "linear", "log_interval": 100}}


def _g(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


# ------------------------------------------------------------------ scheduled-sampling mix (bit-exact: select + multiply)
def test_mix_matches_reference_golden(golden_dir):
    from transformertts_b200.util import apply_teacher_forcing, block_mask
    g = _g(golden_dir, "train_aux")
    pred, mel = torch.from_numpy(g["pred"]).to(DEV), torch.from_numpy(g["mel"]).to(DEV)
    lens, seed = torch.from_numpy(g["lens"]).to(DEV), torch.from_numpy(g["seed_mask"])
    out = apply_teacher_forcing(pred, mel, lens, float(g["p_tf"]), DEV, seed=seed)
    assert np.array_equal(out.cpu().numpy(), g["mixed"])
    bm = block_mask(mel, float(g["p_tf"]), 8, seed=seed)
    assert bm.shape == (3, 37, 1) and torch.equal(bm[:, :, 0].cpu(), TO.dilate_seed(seed, 8))


@pytest.mark.parametrize("B,T,p_tf,L", [(64, 800, 0.7, 8), (1, 1, 0.0, 8), (5, 13, 1.0, 8), (7, 257, 0.5, 3), (2, 40, 0.9, 1)])
def test_mix_matches_oracle(B, T, p_tf, L):
    from transformertts_b200.util import apply_teacher_forcing
    g = torch.Generator().manual_seed(B * T)
    lens = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    mel, pred = torch.randn(B, T, 80, generator=g), torch.randn(B, T, 80, generator=g)
    seed = TO.draw_seed(B, T, p_tf, generator=g)
    ref = TO.apply_teacher_forcing(pred, mel, lens, seed, L)
    out = apply_teacher_forcing(pred.to(DEV), mel.to(DEV), lens.to(DEV), p_tf, DEV, seed=seed, L_bar=L)
    assert torch.equal(out.cpu(), ref)


def test_mix_random_draw_has_the_requested_rate_and_respects_lengths():
    from transformertts_b200.util import apply_teacher_forcing
    B, T = 16, 400
    lens = torch.full((B,), 300, device=DEV)
    pred, mel = torch.ones(B, T, 80, device=DEV), torch.zeros(B, T, 80, device=DEV)
    gen = torch.Generator(device=DEV).manual_seed(3)
    out = apply_teacher_forcing(pred, mel, lens, 0.9, DEV, generator=gen)
    assert float(out[:, 300:].abs().max()) == 0.0
    frac = float(out[:, :300, 0].mean())          # P(frame takes prediction) = 1 - 0.9^8 = 0.57 for L_bar = 8
    assert 0.5 < frac < 0.64, frac
    same = apply_teacher_forcing(pred, mel, lens, 1.0, DEV)
    assert float(same.abs().max()) == 0.0          # p_tf = 1: pure teacher forcing


# ------------------------------------------------------------------ loss
LOSS_RTOL = 2e-5      # fp32 per-thread partial sums + fp64 tree vs torch's fp32 pairwise mean


def test_loss_matches_reference_golden(golden_dir):
    from transformertts_b200 import TransformerTTSLoss
    g = _g(golden_dir, "train_aux")
    out = {"pred_melspec": torch.from_numpy(g["pred"]).to(DEV), "post_melspec": torch.from_numpy(g["post"]).to(DEV),
           "pred_stop": torch.from_numpy(g["stop"]).to(DEV)}
    l = TransformerTTSLoss(stop_weight=8.0)(out, torch.from_numpy(g["mel"]).to(DEV), torch.from_numpy(g["lens"]).to(DEV))
    got = np.array([float(l[k]) for k in ("total", "pred_mel", "post_mel", "stop")])
    print("loss got", got, "reference", g["loss"])
    np.testing.assert_allclose(got, g["loss"], rtol=LOSS_RTOL)
    assert l["total"].dim() == 0 and l["total"].is_cuda


@pytest.mark.parametrize("B,T,w", [(64, 800, 8.0), (1, 1, 8.0), (3, 129, 1.0), (16, 2000, 4.0)])
def test_loss_matches_oracle(B, T, w):
    from transformertts_b200 import TransformerTTSLoss
    g = torch.Generator().manual_seed(B + T)
    lens = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    mel = torch.randn(B, T, 80, generator=g)
    out = {"pred_melspec": mel + 0.5 * torch.randn(B, T, 80, generator=g),
           "post_melspec": mel + 0.25 * torch.randn(B, T, 80, generator=g),
           "pred_stop": 4 * torch.randn(B, T, generator=g)}
    out["pred_melspec"][:, -1] = 1e3                  # garbage in the padding must not reach the loss ...
    lens = torch.clamp(lens, max=max(1, T - 1)) if T > 1 else lens
    ref = TO.loss({k: v.double() for k, v in out.items()}, mel.double(), lens, w)
    crit = TransformerTTSLoss(stop_weight=w)
    dev_out = {k: v.to(DEV) for k, v in out.items()}
    a = crit(dev_out, mel.to(DEV), lens.to(DEV))
    b = crit(dev_out, mel.to(DEV), lens.to(DEV))
    for k in ref:
        np.testing.assert_allclose(float(a[k]), float(ref[k]), rtol=LOSS_RTOL)
        assert float(a[k]) == float(b[k])             # deterministic reduction order


def test_loss_uses_the_pos_weight_buffer_after_load_state_dict():
    from transformertts_b200 import TransformerTTSLoss
    g = torch.Generator().manual_seed(11)
    B, T = 4, 60
    lens = torch.randint(30, T + 1, (B,), generator=g)
    mel = torch.randn(B, T, 80, generator=g)
    out = {"pred_melspec": mel + 0.5, "post_melspec": mel - 0.25, "pred_stop": 4 * torch.randn(B, T, generator=g)}
    crit = TransformerTTSLoss(stop_weight=8.0).to(DEV)
    crit.load_state_dict({"pos_weight": torch.tensor(2.5)})
    ref = TO.loss({k: v.double() for k, v in out.items()}, mel.double(), lens, 2.5)
    got = crit({k: v.to(DEV) for k, v in out.items()}, mel.to(DEV), lens.to(DEV))
    for k in ref:
        np.testing.assert_allclose(float(got[k]), float(ref[k]), rtol=LOSS_RTOL)


# ------------------------------------------------------------------ bf16 mel input, packed outputs, feeder / fetcher
def _model(sd):
    from transformertts_b200 import TransformerTTS
    return TransformerTTS.from_reference(sd, device=DEV)


def test_bf16_mel_input_equals_f32_input_bit_for_bit():
    sd = O.make_state_dict(81)
    ph, mel, pl, ml = O.make_inputs(3, 40, 130, seed=82)         # make_inputs rounds the mel to bf16 already
    m = _model(sd)
    a = m(ph, mel, pl, ml)
    b = m(ph, mel.to(torch.bfloat16), pl, ml)
    torch.cuda.synchronize()
    for k in ("pred_melspec", "post_melspec", "pred_stop"):
        assert torch.equal(a[k], b[k]), k
    for x, y in zip(a["alignments"], b["alignments"]):
        assert torch.equal(x, y)


@pytest.mark.parametrize("dtype,with_align", [(torch.float32, False), (torch.bfloat16, False), (torch.float32, True)])
def test_feeder_and_fetcher_round_trip(dtype, with_align):
    from transformertts_b200.data import OutputFetcher, PinnedFeeder, collate_fn
    sd = O.make_state_dict(83)
    m = _model(sd)
    g = torch.Generator().manual_seed(84)
    batches = []
    for _ in range(5):
        samples = [{"transcript": "t", "phoneme": torch.randint(1, 100, (int(torch.randint(5, 30, (1,), generator=g)),), generator=g),
                    "melspec": O.bf16_round(torch.randn(int(torch.randint(20, 90, (1,), generator=g)), 80, generator=g))}
                   for _ in range(4)]
        batches.append(collate_fn(samples, mel_dtype=torch.bfloat16 if dtype == torch.bfloat16 else torch.float32))
    fetch = OutputFetcher(DEV, dtype=dtype, with_alignments=with_align)
    got = []
    for b in PinnedFeeder(batches, DEV):
        out = m(b["phoneme"], b["melspec"], b["phoneme_lens"], b["melspec_lens"])
        h = fetch(out)
        fetch.wait()
        got.append({k: (v.clone() if k != "alignments" else [a.clone() for a in v]) for k, v in h.items()})
    assert fetch.d2h_bytes > 0
    for b, h in zip(batches, got):
        ref = m(b["phoneme"], b["melspec"], b["phoneme_lens"], b["melspec_lens"])
        torch.cuda.synchronize()
        for k in ("pred_melspec", "post_melspec", "pred_stop"):
            want = ref[k].cpu()
            if dtype == torch.bfloat16:
                assert torch.equal(h[k].float(), want.to(torch.bfloat16).float()), k
            else:
                assert torch.equal(h[k], want), k
        if with_align:
            for x, y in zip(h["alignments"], ref["alignments"]):
                assert torch.equal(x, y.cpu())


def test_packed_weights_follow_in_place_updates_and_moves():
    sd = O.make_state_dict(85)
    ph, mel, pl, ml = O.make_inputs(2, 20, 50, seed=86)
    m = _model(sd)
    a = m(ph, mel, pl, ml)["pred_melspec"].clone()
    with torch.no_grad():
        m.pe.alpha.fill_(0.1)         # an in-place update (optimizer step, BN edit) must reach the kernels
    b = m(ph, mel, pl, ml)["pred_melspec"].clone()
    sd2 = dict(sd)
    sd2["pe.alpha"] = torch.tensor([0.1])
    c = _model(sd2)(ph, mel, pl, ml)["pred_melspec"]
    assert not torch.equal(a, b) and torch.equal(b, c)
    m2 = m.to(DEV)
    assert m2.device == torch.device(DEV) and torch.equal(m2(ph, mel, pl, ml)["pred_melspec"], c)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process():
    """Function attributes (> 48 KB shared memory opt-in) and SM counts are cached per device (host.cu)."""
    sd = O.make_state_dict(87, stop_bias=-1e4)
    ph, mel, pl, ml = O.make_inputs(2, 30, 140, seed=88)
    outs = []
    for d in ("cuda:0", "cuda:1"):
        from transformertts_b200 import TransformerTTS
        m = TransformerTTS.from_reference(sd, device=d)
        o = m(ph, mel, pl, ml)
        i = m.inference(ph, pl, max_len=12)
        torch.cuda.synchronize(d)
        outs.append((o["post_melspec"].cpu(), i["post_melspec"].cpu()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


# ------------------------------------------------------------------ the harness (lightning_module.py bodies)
def _host_batches(n_batches, B, seed):
    g = torch.Generator().manual_seed(seed)
    from transformertts_b200.data import collate_fn
    out = []
    for _ in range(n_batches):
        samples = [{"transcript": "t", "phoneme": torch.randint(1, 100, (int(torch.randint(8, 24, (1,), generator=g)),), generator=g),
                    "melspec": O.bf16_round(torch.randn(int(torch.randint(30, 70, (1,), generator=g)), 80, generator=g))}
                   for _ in range(B)]
        out.append(collate_fn(samples))
    return out


def test_harness_validation_epoch_matches_oracle_pipeline():
    """validation_step (:102-130) over a few batches + on_validation_epoch_end (:137-157): the mean loss against the
    oracle forward + oracle loss on the same batches, and the example inference against the oracle's AR loop."""
    from transformertts_b200.harness import LightningModule
    sd = O.make_state_dict(91)
    lm = LightningModule(CONFIG, device=DEV)
    lm.model.load_state_dict(sd)
    plotted = {}
    lm.plots = {"plot_alignment_single": lambda al, idx, ep, d: plotted.setdefault("align", al[0][idx].cpu()),
                "plot_mels_single": lambda p, t, ep, d: plotted.setdefault("single", p.cpu())}
    batches = _host_batches(3, 4, seed=92)
    avg, pred = lm.run_validation(batches)
    ref_losses = []
    for b in batches:
        r = O.forward(sd, b["phoneme"], b["melspec"], b["phoneme_lens"], b["melspec_lens"])
        ref_losses.append(float(TO.loss(r, b["melspec"], b["melspec_lens"], 8.0)["total"]))
    ref_avg = sum(ref_losses) / len(ref_losses)
    print("harness val_loss", avg, "oracle", ref_avg)
    assert abs(avg - ref_avg) <= 2e-2 * abs(ref_avg)            # bf16 activations: the forward tolerance, averaged
    assert lm.logged["val_loss"] == avg and lm.valid_losses == []
    assert plotted["align"].shape == (4, batches[0]["melspec"].shape[1], batches[0]["phoneme"].shape[1])
    # example inference: phoneme[:1] of the first batch, default max_len / threshold (model/model.py:323)
    assert pred is not None and pred.shape[0] == 1 and pred.shape[2] == 80 and torch.equal(plotted["single"], pred[0].cpu())


def test_harness_training_step_forward_matches_oracle_pipeline():
    """training_step (:45-86) minus the backward: no-grad pass, scheduled-sampling mix, second forward, loss --
    against the oracle running the same sequence with the same seed mask (eval-mode semantics on both sides)."""
    from transformertts_b200.harness import LightningModule
    sd = O.make_state_dict(93)
    lm = LightningModule(CONFIG, device=DEV)
    lm.model.load_state_dict(sd)
    lm.current_epoch = 150                                        # linear schedule: p_tf = 1 - 141/290
    b = _host_batches(1, 5, seed=94)[0]
    B, T = b["melspec"].shape[:2]
    p_tf = TO.get_teacher_forcing_ratio(151, 300, "linear", cycles=1)
    seed = TO.draw_seed(B, T, p_tf, generator=torch.Generator().manual_seed(95))
    total, mixed = lm.training_step_forward(b, 0, seed=seed)
    r1 = O.forward(sd, b["phoneme"], b["melspec"], b["phoneme_lens"], b["melspec_lens"])
    ref_mixed = TO.apply_teacher_forcing(r1["pred_melspec"], b["melspec"], b["melspec_lens"], seed)
    taken = TO.dilate_seed(seed, 8)
    # frames that keep the target are exact; frames that take the prediction carry the forward tolerance
    assert torch.equal(mixed.cpu()[~taken], ref_mixed[~taken])
    assert float((mixed.cpu() - ref_mixed).abs().max()) <= 0.15
    r2 = O.forward(sd, b["phoneme"], ref_mixed, b["phoneme_lens"], b["melspec_lens"])
    ref_total = float(TO.loss(r2, b["melspec"], b["melspec_lens"], 8.0)["total"])
    print("harness train loss", float(total), "oracle", ref_total, "p_tf", p_tf)
    assert abs(float(total) - ref_total) <= 3e-2 * abs(ref_total)
    assert lm.train_losses == [float(total)]
