"""TEST / BASELINE INFRASTRUCTURE ONLY - imports the *unmodified* reference.

Source tree: /root/reference in the build container; on the GPU box (no /root/reference) the byte-for-byte staged
copy oracle/_ref/ written by oracle/stage_ref.py at build() time (git-ignored, sha256 manifest).  It is used
(a) to pin oracle/visinger_oracle.py against the real reference forward,
(b) by oracle/make_golden.py to generate tests/golden/*.npz, and
(c) by bench.py's `cpu_baseline` / `--impl reference` legs as the thing timed on the host cores.

Recipe follows SURVEY.md Appendix A: stub the import-time-only side imports
(librosa, parselmouth, the cp38 monotonic_align .so) and neutralise the hard-coded
`self.text_encoder.cuda(0)` (VISinger/models_ae.py:515, VISinger/models.py:514) while the CPU model is built.
"""
import contextlib
import os
import sys
import types

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
STAGED_ROOT = os.path.join(_HERE, "_ref")


def _resolve_root():
    env = os.environ.get("B200SVS_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/VISinger"):
        return "/root/reference"
    return STAGED_ROOT


REF_ROOT = _resolve_root()


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "VISinger", "models_ae.py"))


def source_kind() -> str:
    """'tree' (the read-only reference checkout) or 'staged' (oracle/_ref, the byte-for-byte copy)."""
    return "staged" if os.path.abspath(REF_ROOT) == os.path.abspath(STAGED_ROOT) else "tree"


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


_imported = {}


@contextlib.contextmanager
def _cpu_only_cuda():
    """`Module.cuda()` is the identity while a CPU reference model is constructed (models_ae.py:515)."""
    orig = torch.nn.Module.cuda
    torch.nn.Module.cuda = lambda self, *a, **k: self
    try:
        yield
    finally:
        torch.nn.Module.cuda = orig


def import_reference():
    """Returns (models_ae, models, vutils) reference modules."""
    if _imported:
        return _imported["ae"], _imported["v4"], _imported["utils"]
    if not available():
        raise RuntimeError("reference tree not present at %s (run oracle/stage_ref.py in the build container)" % REF_ROOT)
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    if "librosa" not in sys.modules:
        f = _stub("librosa.filters", mel=lambda *a, **k: None)
        _stub("librosa", filters=f)
    if "parselmouth" not in sys.modules:
        _stub("parselmouth")
    _stub("VISinger.monotonic_align", maximum_path=None)
    import logging
    lvl = logging.getLogger().level
    from VISinger import models_ae, models
    from VISinger.utils import utils as vutils
    logging.getLogger().setLevel(max(lvl, logging.WARNING))
    _imported.update(ae=models_ae, v4=models, utils=vutils)
    return models_ae, models, vutils


def build_reference(variant="ae", gin_channels=0, seed=1234):
    """variant: 'ae' (v5/v6, VISinger/models_ae.py) or 'v4' (VISinger/models.py).  Always a CPU model."""
    models_ae, models, vutils = import_reference()
    cfg = "ours_ae.json" if variant == "ae" else "ours.json"
    hps = vutils.get_hparams_from_file(os.path.join(REF_ROOT, "VISinger/configs", cfg))
    mod = models_ae if variant == "ae" else models
    torch.manual_seed(seed)
    kw = dict(hps.model)
    if gin_channels:
        kw["gin_channels"] = gin_channels
    with _cpu_only_cuda():
        net = mod.SynthesizerTrn(100, hps.data.filter_length // 2 + 1,
                                 hps.train.segment_size // hps.data.hop_length, **kw).eval()
    return net


def enlarge_positions(net, n):
    """The reference's position table has 2000 rows (ext_modules/text_post_net.py:26,31-32); T > 2000 configs
     replace the two buffers by larger ones before the state-dict with the enlarged table is loaded.
    The forward code is untouched."""
    tp = net.text_post_net
    if tp.position_embeddings.weight.shape[0] >= n:
        return net
    tp.position_embeddings = torch.nn.Embedding(n, tp.position_embeddings.weight.shape[1])
    tp.position_ids = torch.arange(n).expand((1, -1))
    return net


def run_reference_infer(net, inputs, noise, ph_dur=None, **kw):
    """Runs net.infer with torch.randn_like patched to return `noise` (models_ae.py:643)."""
    orig = torch.randn_like
    torch.randn_like = lambda t, *a, **k: noise.to(t.dtype)
    try:
        with torch.no_grad():
            return net.infer(inputs["x"], inputs["x_lengths"], inputs["note_pitch_id"],
                             inputs["note_duration"], phone_shengdiao=inputs["phone_shengdiao"],
                             ph_dur=ph_dur, noise_scale=1., length_scale=1, **kw)
    finally:
        torch.randn_like = orig
