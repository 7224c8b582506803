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


def _reference_util_functions():
    """utils/util.py imports matplotlib and loguru at module level (absent here), so the reference's OWN source of the
    three functions on the path is cut out with `ast` and executed unmodified in a namespace that holds what they use."""
    import ast
    import math
    import torch.nn.functional as F
    src = open("/root/reference/utils/util.py").read()
    want = {"block_mask", "apply_teacher_forcing", "get_teacher_forcing_ratio"}
    ns = {"torch": torch, "F": F, "math": math, "Tensor": torch.Tensor}
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in want:
            exec(compile(ast.Module(body=[node], type_ignores=[]), "/root/reference/utils/util.py", "exec"), ns)
    assert want <= set(ns)
    return ns


def train_aux_case(name, seed, B, T, p_tf):
    """Scheduled-sampling mix + loss: the reference's own code (loss.py imported, utils/util.py functions executed from
    source) against oracle/train_oracle.py; inputs and reference outputs go into the fixture."""
    import train_oracle as TO
    from loss import TransformerTTSLoss
    ns = _reference_util_functions()
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(T // 2, T + 1, (B,), generator=g)
    lens[0] = T
    mel = torch.randn(B, T, 80, generator=g) * (torch.arange(T)[None, :] < lens[:, None])[:, :, None]
    pred = mel + 0.3 * torch.randn(B, T, 80, generator=g)
    post = mel + 0.2 * torch.randn(B, T, 80, generator=g)
    stop = torch.randn(B, T, generator=g) * 3
    # the reference draws its seed mask from the global torch RNG: replay the same draw for the oracle
    torch.manual_seed(seed + 7)
    ref_mixed = ns["apply_teacher_forcing"](pred, mel, lens, p_tf, "cpu")
    torch.manual_seed(seed + 7)
    seed_mask = TO.draw_seed(B, T, p_tf)
    ora_mixed = TO.apply_teacher_forcing(pred, mel, lens, seed_mask)
    assert torch.equal(ref_mixed, ora_mixed), "scheduled-sampling mix: oracle differs from the reference"
    frac = float(TO.dilate_seed(seed_mask, 8).float().mean())
    ref_loss = TransformerTTSLoss(stop_weight=8.0)({"pred_melspec": pred, "post_melspec": post, "pred_stop": stop}, mel, lens)
    ora_loss = TO.loss({"pred_melspec": pred, "post_melspec": post, "pred_stop": stop}, mel, lens, 8.0)
    for k in ref_loss:
        d = abs(float(ref_loss[k]) - float(ora_loss[k]))
        print(name, "loss", k, float(ref_loss[k]), "oracle diff", d)
        assert d <= 1e-6 * max(1.0, abs(float(ref_loss[k])))
    sched = [(e, m) for m in ("cosine", "linear", "constant") for e in (1, 9, 10, 11, 100, 155, 299, 300, 400)]
    ref_sched = [ns["get_teacher_forcing_ratio"](epoch=e, total_epochs=300, mode=m, cycles=1) for e, m in sched]
    ora_sched = [TO.get_teacher_forcing_ratio(epoch=e, total_epochs=300, mode=m, cycles=1) for e, m in sched]
    assert ref_sched == ora_sched
    print(name, "masked fraction", frac, "p_tf", p_tf)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), seed=seed, p_tf=p_tf, lens=lens.numpy(), mel=mel.numpy(),
                        pred=pred.numpy(), post=post.numpy(), stop=stop.numpy(), seed_mask=seed_mask.numpy(),
                        mixed=ref_mixed.numpy(),
                        loss=np.array([float(ref_loss[k]) for k in ("total", "pred_mel", "post_mel", "stop")], dtype=np.float64),
                        sched_epochs=np.array([e for e, _ in sched]), sched_modes=np.array([m for _, m in sched]),
                        sched=np.array(ref_sched, dtype=np.float64))


def collate_case(name, seed, B):
    """dataset.py:71-102 `collate_fn`, executed from the reference's own source (the module imports pytorch_lightning,
    absent here), on a seeded ragged list of samples; the samples and the collated batch go into the fixture."""
    import ast
    src = open("/root/reference/dataset.py").read()
    ns = {"torch": torch}
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == "collate_fn":
            exec(compile(ast.Module(body=[node], type_ignores=[]), "/root/reference/dataset.py", "exec"), ns)
    g = torch.Generator().manual_seed(seed)
    lp = torch.randint(3, 20, (B,), generator=g)
    lp[1] = lp[3]                                            # a tie in the sort key
    lm = torch.randint(5, 40, (B,), generator=g)
    samples = [{"transcript": f"utt{i}", "phoneme": torch.randint(1, 100, (int(lp[i]),), generator=g),
                "melspec": torch.randn(int(lm[i]), 80, generator=g)} for i in range(B)]
    out = ns["collate_fn"](samples)
    rec = {"n": B}
    for i, smp in enumerate(samples):
        rec[f"ph{i}"], rec[f"mel{i}"] = smp["phoneme"].numpy(), smp["melspec"].numpy()
    for k in ("phoneme", "melspec", "phoneme_lens", "melspec_lens"):
        rec["out_" + k] = out[k].numpy()
    rec["out_transcript"] = np.array(out["transcript"])
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(name, "collated", tuple(out["phoneme"].shape), tuple(out["melspec"].shape), out["transcript"])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    if "--only-train-aux" in sys.argv:
        train_aux_case("train_aux", 61, B=3, T=37, p_tf=0.8)
        collate_case("collate", 62, B=5)
        sys.exit(0)
    fwd_case("fwd_small", 11, B=3, Tp=12, Tm=40)
    fwd_case("fwd_tile", 12, B=2, Tp=37, Tm=150, keep_align=False)     # crosses a 128-row tile
    fwd_case("fwd_b1", 13, B=1, Tp=9, Tm=17)
    inf_case("inf_b1", 44, B=1, Tp=14, horizon=48, min_t=10)
    inf_case("inf_b3", 38, B=3, Tp=11, horizon=48, min_t=10)         # ragged: encoder unmasked (F8)
    train_aux_case("train_aux", 61, B=3, T=37, p_tf=0.8)
    collate_case("collate", 62, B=5)
    print("golden fixtures written to", os.path.normpath(OUT))
