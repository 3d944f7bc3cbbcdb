"""Drop-in for `VISinger.models_ae` (SVS v5 / v6): `from music_generation_b200 import models_ae`."""
from .synthesizer import SynthesizerTrn  # noqa: F401  (variant "ae", 6-tuple infer result)
