"""TEST INFRASTRUCTURE ONLY - imports the *unmodified* reference from /root/reference.

Usable only in the build container (the GPU box has no /root/reference).  It is used
(a) to pin oracle/visinger_oracle.py against the real reference forward and
(b) by oracle/make_golden.py to generate tests/golden/*.npz.

Recipe follows SURVEY.md Appendix A: stub the import-time-only side imports
(librosa, parselmouth, the cp38 monotonic_align .so) and neutralise the hard-coded
`self.text_encoder.cuda(0)` (VISinger/models_ae.py:515, VISinger/models.py:514).
"""
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("B200SVS_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "VISinger"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


_imported = {}


def import_reference():
    """Returns (models_ae, models, vutils) reference modules."""
    if _imported:
        return _imported["ae"], _imported["v4"], _imported["utils"]
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    f = _stub("librosa.filters", mel=lambda *a, **k: None)
    _stub("librosa", filters=f)
    _stub("parselmouth")
    _stub("VISinger.monotonic_align", maximum_path=None)
    if not torch.cuda.is_available():
        torch.nn.Module.cuda = lambda self, *a, **k: self
    import logging
    lvl = logging.getLogger().level
    from VISinger import models_ae, models
    from VISinger.utils import utils as vutils
    logging.getLogger().setLevel(max(lvl, logging.WARNING))
    _imported.update(ae=models_ae, v4=models, utils=vutils)
    return models_ae, models, vutils


def build_reference(variant="ae", gin_channels=0, seed=1234):
    """variant: 'ae' (v5/v6, VISinger/models_ae.py) or 'v4' (VISinger/models.py)."""
    models_ae, models, vutils = import_reference()
    cfg = "ours_ae.json" if variant == "ae" else "ours.json"
    hps = vutils.get_hparams_from_file(os.path.join(REF_ROOT, "VISinger/configs", cfg))
    mod = models_ae if variant == "ae" else models
    torch.manual_seed(seed)
    kw = dict(hps.model)
    if gin_channels:
        kw["gin_channels"] = gin_channels
    net = mod.SynthesizerTrn(100, hps.data.filter_length // 2 + 1,
                             hps.train.segment_size // hps.data.hop_length, **kw).eval()
    return net


def run_reference_infer(net, inputs, noise, ph_dur=None):
    """Runs net.infer with torch.randn_like patched to return `noise` (models_ae.py:643)."""
    orig = torch.randn_like
    torch.randn_like = lambda t, *a, **k: noise.to(t.dtype)
    try:
        with torch.no_grad():
            return net.infer(inputs["x"], inputs["x_lengths"], inputs["note_pitch_id"],
                             inputs["note_duration"], phone_shengdiao=inputs["phone_shengdiao"],
                             ph_dur=ph_dur, noise_scale=1., length_scale=1)
    finally:
        torch.randn_like = orig
