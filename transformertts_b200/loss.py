"""Drop-in for the reference's `loss.TransformerTTSLoss` (loss.py:9-55) on B200, forward only.

Same constructor (`stop_weight`), same call `criterion(outputs, mel, lengths)` and the same dict of four scalar
tensors ("total", "pred_mel", "post_mel", "stop").  The three masked reductions run in one hand-written streaming
kernel over the forward's outputs in place (csrc/train_aux.cu through `tts_loss`); padded frames are never read.
There is no autograd: the backward pass is SURVEY 8(f) rank 4.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict

import torch
import torch.nn as nn

from . import _ext

N_MELS = 80


class TransformerTTSLoss(nn.Module):
    def __init__(self, stop_weight: float = 8.0):
        super().__init__()
        self.register_buffer("pos_weight", torch.tensor(float(stop_weight)))      # loss.py:13
        self._stop_weight = float(stop_weight)  # host copy of the buffer, refreshed when the buffer changes (_weight())
        self._pw_seen = (self.pos_weight, self.pos_weight._version)
        self._scratch = None
        self.values = None                     # [total, pred_mel, post_mel, stop] of the last call (device tensor)

    def _weight(self) -> float:
        """pos_weight as the reference uses it (the registered buffer, loss.py:13,44): a value restored by
        load_state_dict, replaced by .to() or edited in place is picked up; the device -> host read happens only then."""
        pw = self.pos_weight
        if self._pw_seen[0] is not pw or self._pw_seen[1] != pw._version:
            self._stop_weight = float(pw)
            self._pw_seen = (pw, pw._version)
        return self._stop_weight

    def forward(self, outputs: Dict[str, torch.Tensor], mel: torch.Tensor, lengths: torch.Tensor) -> Dict[str, torch.Tensor]:
        lib = _ext.load()
        pred, post, stop = outputs["pred_melspec"], outputs["post_melspec"], outputs["pred_stop"]
        if not pred.is_cuda:
            raise RuntimeError("transformertts_b200.TransformerTTSLoss runs on the GPU only (no CPU path)")
        dev = pred.device
        B, T, Cm = pred.shape
        if Cm != N_MELS or post.shape != pred.shape or stop.shape != (B, T) or mel.shape != pred.shape:
            raise ValueError("expected pred/post/mel [B,T,80] and pred_stop [B,T]")
        f32 = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()
        pred, post, stop, mel = f32(pred), f32(post), f32(stop), f32(mel)
        lengths = lengths.to(device=dev, dtype=torch.int64).contiguous()
        with torch.cuda.device(dev):
            need = lib.tts_loss_scratch_bytes()
            if self._scratch is None or self._scratch.device != dev or self._scratch.numel() * 8 < need:
                self._scratch = torch.empty((need + 7) // 8, dtype=torch.float64, device=dev)
            out = torch.empty(4, dtype=torch.float32, device=dev)
            _ext.check(lib.tts_loss(_ext.ptr(pred), _ext.ptr(post), _ext.ptr(stop), _ext.ptr(mel), _ext.ptr(lengths), B, T,
                                    self._weight(), _ext.ptr(out), _ext.ptr(self._scratch), self._scratch.numel() * 8,
                                    C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)), "tts_loss")
        self.values = out                      # the four scalars as one tensor: one copy brings them all to the host
        return {"total": out[0], "pred_mel": out[1], "post_mel": out[2], "stop": out[3]}   # loss.py:49-54
