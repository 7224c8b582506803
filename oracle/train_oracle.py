"""CPU oracle for the forward-only pieces of the reference's training / validation step  --  TEST INFRASTRUCTURE,
NOT PRODUCT (only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it).

Restates, in plain torch on the CPU:
  * block_mask / apply_teacher_forcing   utils/util.py:103-120   (block-wise scheduled sampling, L_bar = 8)
  * TransformerTTSLoss.forward           loss.py:15-55           (masked MSE x2 + BCE-with-logits, pos_weight 8)
  * get_teacher_forcing_ratio            utils/util.py:52-92     (epoch schedule; host arithmetic)

Parity pin: `oracle/make_golden.py` runs the reference's own code for these functions in the build container -- the
`loss.TransformerTTSLoss` class imported from /root/reference, and the source text of the three utils/util.py functions
extracted with `ast` and executed as-is (the module itself imports matplotlib / loguru, which are absent) -- and commits
inputs + outputs as tests/golden/train_aux.npz; tests/test_oracle.py re-checks this file against that fixture.
The random seed mask of block_mask (`torch.rand(B,1,T) < 1 - p_tf`) is an INPUT here so the comparison is deterministic.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


def dilate_seed(seed: Tensor, L_bar: int) -> Tensor:
    """utils/util.py:108-110: max_pool1d(seed, kernel_size=L_bar, stride=1, padding=L_bar//2), cut back to T.
    seed: (B,T) bool/0-1  ->  (B,T) bool."""
    B, T = seed.shape
    d = F.max_pool1d(seed.float().unsqueeze(1), kernel_size=L_bar, stride=1, padding=L_bar // 2)
    return d.squeeze(1).bool()[:, :T]


def draw_seed(B: int, T: int, p_tf: float, generator=None, device="cpu") -> Tensor:
    """utils/util.py:108: seed = torch.rand(B, 1, T) < (1 - p_tf)   ->  (B,T) bool."""
    return (torch.rand(B, 1, T, generator=generator, device=device) < (1 - p_tf)).squeeze(1)


def apply_teacher_forcing(pred_melspec: Tensor, melspec: Tensor, melspec_lens: Tensor, seed: Tensor,
                          L_bar: int = 8) -> Tensor:
    """utils/util.py:113-120 with the random seed mask passed in."""
    mask = dilate_seed(seed, L_bar).unsqueeze(-1)                                   # :117 (block_mask)
    mel_mixed = torch.where(mask, pred_melspec, melspec)                            # :118
    valid = torch.arange(pred_melspec.size(1)).unsqueeze(0) < melspec_lens.unsqueeze(1)   # :119
    return mel_mixed * valid.unsqueeze(-1)                                          # :120


def loss(outputs: Dict[str, Tensor], mel: Tensor, lengths: Tensor, stop_weight: float = 8.0) -> Dict[str, Tensor]:
    """loss.py:15-55."""
    pred, post, stop = outputs["pred_melspec"], outputs["post_melspec"], outputs["pred_stop"]
    B, T, C = pred.shape
    rng = torch.arange(T).unsqueeze(0).expand(B, T)
    mask = rng < lengths.unsqueeze(1)                                               # :28-29
    gate = rng == (lengths.unsqueeze(1) - 1)                                        # :30
    true = mel[mask].view(-1, C)
    pred_mel = ((pred[mask].view(-1, C) - true) ** 2).mean()                        # :37  nn.MSELoss()
    post_mel = ((post[mask].view(-1, C) - true) ** 2).mean()                        # :38
    x, y = stop[mask], gate.float()[mask]
    sp = torch.log1p(torch.exp(-x.abs())) + torch.clamp(-x, min=0)                  # softplus(-x)
    stop_loss = ((1 - y) * x + (1 + (stop_weight - 1) * y) * sp).mean()             # :42-46  BCE-with-logits, pos_weight
    return {"total": pred_mel + 0.5 * post_mel + stop_loss, "pred_mel": pred_mel, "post_mel": post_mel,
            "stop": stop_loss}                                                      # :39,48-54


def get_teacher_forcing_ratio(epoch: int, total_epochs: int = 300, mode: str = "cosine", warmup_epochs: int = 10,
                              **kwargs) -> float:
    """utils/util.py:52-92."""
    if epoch < warmup_epochs:
        return 1.0
    e, tot = epoch - warmup_epochs, max(total_epochs - warmup_epochs, 1)
    if mode == "cosine":
        r = 0.5 * math.cos(math.pi * e * kwargs.get("cycles", 1) / tot) + 0.5
        return max(min(r, 1.0), 0.5)
    if mode == "linear":
        return max(1.0 - e / tot, 0.05)
    if mode == "constant":
        return kwargs.get("value", 1.0)
    raise ValueError(f"Unsupported teacher forcing mode: {mode}")
