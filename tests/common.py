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
