"""Drop-in for `VISinger.models` (SVS v1-v4): no spec_decoder / spec_proj, 5-tuple infer result."""
from .synthesizer import SynthesizerTrn as _Base


class SynthesizerTrn(_Base):
    variant = "v4"
