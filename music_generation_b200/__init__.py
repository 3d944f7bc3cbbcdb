"""music_generation_b200 - B200-native VISinger synthesis forward (drop-in for SynthesizerTrn.infer)."""
from . import models, models_ae  # noqa: F401
from ._lib import build  # noqa: F401

__all__ = ["models", "models_ae", "build"]
