"""CPU tests (no GPU): the train/validation oracle (oracle/train_oracle.py) against the golden vectors produced by the
reference's own code (oracle/make_golden.py: loss.py imported, utils/util.py and dataset.py functions executed from
source), and the host logic of transformertts_b200.data / .util against the same fixtures."""
import os

import numpy as np
import pytest
import torch

import train_oracle as TO


def _g(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def test_mix_oracle_matches_reference_golden(golden_dir):
    g = _g(golden_dir, "train_aux")
    mixed = TO.apply_teacher_forcing(torch.from_numpy(g["pred"]), torch.from_numpy(g["mel"]), torch.from_numpy(g["lens"]),
                                     torch.from_numpy(g["seed_mask"]))
    assert np.array_equal(mixed.numpy(), g["mixed"])                 # select + multiply: bit-exact


def test_dilation_is_window_minus4_plus3():
    """max_pool1d(kernel 8, stride 1, padding 4)[:T]: frame t takes the prediction iff a seed lies in [t-4, t+3]."""
    seed = torch.zeros(1, 30, dtype=torch.bool)
    seed[0, 10] = True
    d = TO.dilate_seed(seed, 8)[0]
    assert d.nonzero().flatten().tolist() == list(range(7, 15))
    seed = torch.zeros(1, 30, dtype=torch.bool)
    seed[0, 0] = seed[0, 29] = True
    d = TO.dilate_seed(seed, 8)[0]
    assert d.nonzero().flatten().tolist() == [0, 1, 2, 3, 4, 26, 27, 28, 29]


def test_loss_oracle_matches_reference_golden(golden_dir):
    g = _g(golden_dir, "train_aux")
    out = {"pred_melspec": torch.from_numpy(g["pred"]), "post_melspec": torch.from_numpy(g["post"]),
           "pred_stop": torch.from_numpy(g["stop"])}
    l = TO.loss(out, torch.from_numpy(g["mel"]), torch.from_numpy(g["lens"]), 8.0)
    got = np.array([float(l[k]) for k in ("total", "pred_mel", "post_mel", "stop")])
    np.testing.assert_allclose(got, g["loss"], rtol=1e-6)


def test_teacher_forcing_schedule_matches_reference_golden(golden_dir):
    from transformertts_b200.util import get_teacher_forcing_ratio
    g = _g(golden_dir, "train_aux")
    for e, m, want in zip(g["sched_epochs"], g["sched_modes"], g["sched"]):
        assert TO.get_teacher_forcing_ratio(epoch=int(e), total_epochs=300, mode=str(m), cycles=1) == want
        assert get_teacher_forcing_ratio(epoch=int(e), total_epochs=300, mode=str(m), cycles=1) == want
    with pytest.raises(ValueError):
        get_teacher_forcing_ratio(epoch=50, mode="bogus")


@pytest.mark.parametrize("mel_dtype", [torch.float32, torch.bfloat16])
def test_collate_matches_reference_golden(golden_dir, mel_dtype):
    from transformertts_b200.data import collate_fn
    g = _g(golden_dir, "collate")
    samples = [{"transcript": f"utt{i}", "phoneme": torch.from_numpy(g[f"ph{i}"]), "melspec": torch.from_numpy(g[f"mel{i}"])}
               for i in range(int(g["n"]))]
    out = collate_fn(samples, mel_dtype=mel_dtype)
    assert np.array_equal(out["phoneme"].numpy(), g["out_phoneme"])
    assert np.array_equal(out["phoneme_lens"].numpy(), g["out_phoneme_lens"])
    assert np.array_equal(out["melspec_lens"].numpy(), g["out_melspec_lens"])
    assert out["transcript"] == [str(t) for t in g["out_transcript"]]
    want = torch.from_numpy(g["out_melspec"])
    if mel_dtype == torch.float32:
        assert torch.equal(out["melspec"], want)
    else:
        assert out["melspec"].dtype == torch.bfloat16 and torch.equal(out["melspec"], want.to(torch.bfloat16))
    assert out["phoneme"].dtype == torch.long and int(out["phoneme_lens"][0]) == out["phoneme"].shape[1]


def test_bucket_batches_cover_every_utterance_once_and_cut_padding():
    from transformertts_b200.data import bucket_batches, padding_efficiency
    g = torch.Generator().manual_seed(5)
    lens = torch.randint(200, 870, (1003,), generator=g).tolist()          # LJSpeech-shaped mel lengths
    plain = [list(range(s, min(s + 16, len(lens)))) for s in range(0, len(lens), 16)]
    for shuffle in (False, True):
        b = bucket_batches(lens, 16, shuffle=shuffle, generator=torch.Generator().manual_seed(1))
        flat = sorted(i for bb in b for i in bb)
        assert flat == list(range(len(lens)))
        assert all(len(bb) <= 16 for bb in b) and sum(len(bb) < 16 for bb in b) <= (21 if shuffle else 1)
        assert padding_efficiency(lens, b) > 0.95 > padding_efficiency(lens, plain)
    d = bucket_batches(lens, 16, drop_last=True)
    assert all(len(bb) == 16 for bb in d)
    assert bucket_batches([], 4) == []


def test_product_modules_do_not_import_the_oracle_or_the_staged_reference():
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "transformertts_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(import|from)\s+(tts_oracle|train_oracle|make_ref|oracle)\b", src, re.M), fn
            assert "oracle/_ref" not in src and "oracle._ref" not in src, fn


def test_gpu_only_entry_points_fail_loudly_on_cpu_tensors():
    from transformertts_b200 import TransformerTTSLoss
    from transformertts_b200.util import apply_teacher_forcing
    x = torch.zeros(1, 4, 80)
    with pytest.raises(RuntimeError, match="GPU only"):
        TransformerTTSLoss()({"pred_melspec": x, "post_melspec": x, "pred_stop": torch.zeros(1, 4)}, x, torch.tensor([4]))
    with pytest.raises(RuntimeError, match="GPU only"):
        apply_teacher_forcing(x, x, torch.tensor([4]), 0.5)


def test_loss_pos_weight_follows_the_registered_buffer():
    """loss.py:13,44: the reference reads the `pos_weight` buffer, so a checkpointed or edited value must win over the
    constructor argument."""
    from transformertts_b200 import TransformerTTSLoss
    crit = TransformerTTSLoss(stop_weight=8.0)
    assert crit._weight() == 8.0
    crit.load_state_dict({"pos_weight": torch.tensor(3.0)})
    assert crit._weight() == 3.0
    crit.pos_weight.fill_(5.0)
    assert crit._weight() == 5.0
    crit.pos_weight = torch.tensor(2.0)
    assert crit._weight() == 2.0
