"""B200-native TransformerTTS hot path: teacher-forced forward and KV-cached AR decode.

`TransformerTTS` mirrors the reference's `model.TransformerTTS` (model/model.py:138-394); the compute is
hand-written CUDA for sm_100a behind the C ABI of include/tts_b200.h.
"""
from .api import TransformerTTS  # noqa: F401
from .loss import TransformerTTSLoss  # noqa: F401
from . import _ext  # noqa: F401

__all__ = ["TransformerTTS", "TransformerTTSLoss"]
