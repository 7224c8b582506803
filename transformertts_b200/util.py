"""B200 counterparts of the reference's utils/util.py functions that sit on the forward-only training / validation path:
`prepare_batch` (:96-100), `block_mask` (:103-110), `apply_teacher_forcing` (:113-120) and the host-side schedule
`get_teacher_forcing_ratio` (:52-92).  Same names, argument meaning and results; the mixing itself is one CUDA kernel
(csrc/train_aux.cu, `tts_teacher_forcing_mix`).  The logging / plotting / LR-schedule helpers of that file are control
plane and out of scope.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, Optional, Tuple

import torch

from . import _ext

Tensor = torch.Tensor


def prepare_batch(batch: Dict[str, Tensor], device) -> Tuple[Tensor, ...]:
    """utils/util.py:96-100: move a collated batch to the target device (non-blocking: pinned batches overlap)."""
    return tuple(batch[k].to(device, non_blocking=True) for k in ["phoneme", "melspec", "phoneme_lens", "melspec_lens"])


def get_teacher_forcing_ratio(epoch: int, total_epochs: int = 300, mode: str = "cosine", warmup_epochs: int = 10,
                              **kwargs) -> float:
    """utils/util.py:52-92: 1.0 during warm-up, then cosine (floor 0.5), linear (floor 0.05) or constant."""
    if epoch < warmup_epochs:
        return 1.0
    e, tot = epoch - warmup_epochs, max(total_epochs - warmup_epochs, 1)
    if mode == "cosine":
        ratio = 0.5 * math.cos(math.pi * e * kwargs.get("cycles", 1) / tot) + 0.5
        return max(min(ratio, 1.0), 0.5)
    if mode == "linear":
        return max(1.0 - e / tot, 0.05)
    if mode == "constant":
        return kwargs.get("value", 1.0)
    raise ValueError(f"Unsupported teacher forcing mode: {mode}")


def draw_seed(mel: Tensor, p_tf: float, generator: Optional[torch.Generator] = None) -> Tensor:
    """The random half of block_mask (utils/util.py:108): `torch.rand(B,1,T) < 1 - p_tf` on mel's device -> (B,T) uint8."""
    B, T, _ = mel.shape
    return (torch.rand(B, 1, T, device=mel.device, generator=generator) < (1 - p_tf)).squeeze(1).to(torch.uint8)


def apply_teacher_forcing(pred_melspec: Tensor, melspec: Tensor, melspec_lens: Tensor, p_tf: float, device=None,
                          seed: Optional[Tensor] = None, generator: Optional[torch.Generator] = None,
                          L_bar: int = 8) -> Tensor:
    """utils/util.py:113-120: block-wise scheduled sampling.  `seed` (B,T) 0/1 replaces the random draw when given
    (deterministic parity tests); otherwise it is drawn with torch's CUDA generator exactly as the reference does."""
    lib = _ext.load()
    if not pred_melspec.is_cuda:
        raise RuntimeError("transformertts_b200.util.apply_teacher_forcing runs on the GPU only (no CPU path)")
    dev = pred_melspec.device
    B, T, Cm = pred_melspec.shape
    if Cm != 80 or melspec.shape != pred_melspec.shape:
        raise ValueError("expected pred_melspec and melspec [B,T,80]")
    pred = pred_melspec.detach().to(torch.float32).contiguous()
    mel = melspec.to(device=dev, dtype=torch.float32).contiguous()
    lens = melspec_lens.to(device=dev, dtype=torch.int64).contiguous()
    if seed is None:
        seed = draw_seed(pred, p_tf, generator)
    seed = seed.to(device=dev, dtype=torch.uint8).contiguous()
    if seed.shape != (B, T):
        raise ValueError("seed must be [B,T]")
    out = torch.empty_like(pred)
    with torch.cuda.device(dev):
        _ext.check(lib.tts_teacher_forcing_mix(_ext.ptr(pred), _ext.ptr(mel), _ext.ptr(seed), _ext.ptr(lens), B, T,
                                               int(L_bar), _ext.ptr(out),
                                               C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)),
                   "tts_teacher_forcing_mix")
    return out


def block_mask(mel: Tensor, p_tf: float, L_bar: int, seed: Optional[Tensor] = None,
               generator: Optional[torch.Generator] = None) -> Tensor:
    """utils/util.py:103-110: (B,T,1) bool mask of the frames that take the model's own prediction.  Computed with the
    mix kernel itself (ones vs zeros through the same dilation), so it is exactly the mask `apply_teacher_forcing` uses."""
    B, T, _ = mel.shape
    if seed is None:
        seed = draw_seed(mel, p_tf, generator)
    ones = torch.ones(B, T, 80, device=mel.device)
    full = torch.full((B,), T, dtype=torch.int64, device=mel.device)
    mixed = apply_teacher_forcing(ones, torch.zeros_like(ones), full, p_tf, seed=seed, L_bar=L_bar)
    return mixed[:, :, :1] > 0.5
