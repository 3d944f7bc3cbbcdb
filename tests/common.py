"""Shared helpers of the test-suite (metrics, fixture loading, deterministic weights)."""
import functools
import os

import numpy as np
import torch

from music_generation_b200 import models, models_ae, synthetic
from conftest import HP

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# waveform tolerance stated by BASELINE.json north_star
MAX_ABS_TOL = 2e-3
SNR_DB_TOL = 40.0


def snr_db(ref, out):
    ref, out = ref.double(), out.double()
    return float(10 * torch.log10((ref ** 2).sum() / ((ref - out) ** 2).sum().clamp_min(1e-300)))


def max_abs(ref, out):
    return float((ref.double() - out.double()).abs().max())


@functools.lru_cache(maxsize=4)
def state_dict_for(variant, seed=1234, gin=0):
    """Same recipe as oracle/make_golden.py: seeded package init + synthetic.apply_recipe."""
    mod = models_ae if variant == "ae" else models
    net = mod.SynthesizerTrn(100, 513, 48, **HP, seed=seed, gin_channels=gin)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    return synthetic.apply_recipe(sd, seed=seed)


def build_model(variant, gin=0, device="cpu", **kw):
    """Package model loaded with the deterministic recipe weights."""
    mod = models_ae if variant == "ae" else models
    net = mod.SynthesizerTrn(100, 513, 48, **HP, gin_channels=gin, **kw).eval()
    sd = state_dict_for(variant, gin=gin)
    net.load_state_dict(sd)
    return net.to(device), sd


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


def flat(res):
    return [res[0], res[1], *res[2], *res[3:]]


_REPORT = {}


def report(key, value):
    """Records a parity figure that a test REPORTS rather than asserts to be zero (f0 bucket / duration flips on
    quantisation edges, SNRs): printed (visible with -s / -rA) and accumulated in gpurun_out/parity_report.json."""
    import json
    _REPORT[key] = value
    print(f"[parity] {key} = {value}")
    out = os.path.join(os.path.dirname(GOLDEN), "..", "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        path = os.path.join(out, "parity_report.json")
        old = json.load(open(path)) if os.path.exists(path) else {}
        old.update(_REPORT)
        json.dump(old, open(path, "w"), indent=1, sort_keys=True)
    except OSError:
        pass


def coarse_edge_distance(log_f0):
    """Distance (in coarse-bucket units) of f0_to_coarse's pre-rounding value from the nearest .5 rounding edge
    (utils/pitch_utils.py:22-31): a bucket may only differ between two fp32 evaluations where this is ~0."""
    import math
    f0 = torch.exp(log_f0.double()) - 1.0
    mel = 1127.0 * torch.log1p(f0.clamp_min(0) / 700.0)
    lo, hi = 1127.0 * math.log1p(50.0 / 700.0), 1127.0 * math.log1p(1100.0 / 700.0)
    v = torch.where(mel > 0, (mel - lo) * 254.0 / (hi - lo) + 1.0, mel)
    frac = (v + 0.5) - torch.floor(v + 0.5)
    return torch.minimum(frac, 1.0 - frac)
