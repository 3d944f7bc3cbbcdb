"""-m gpu: the CTA-pair kernel (conv1d_tc_kernel<true>, tcgen05 cta_group::2) against the single-CTA kernel on the same
inputs - the results must be bit-identical (same MMA K order, same epilogue).  B200SVS_TC_PAIR_TILES=2 makes the
library pair even these small launches (production pairs only launches of >= 2 tiles per SM)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _run(tmp, name, env_extra):
    path = os.path.join(tmp, name)
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, os.path.join(HERE, "pair_worker.py"), path], env=env, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return torch.load(path)


def test_cta_pair_kernel_is_bit_identical_to_single_cta(tmp_path):
    single = _run(str(tmp_path), "single.pt", {"B200SVS_TC_PAIR": "0"})
    paired = _run(str(tmp_path), "pair.pt", {"B200SVS_TC_PAIR": "1", "B200SVS_TC_PAIR_TILES": "2", "B200SVS_TC_PAIR_MIN": "1"})
    assert single.keys() == paired.keys() and len(single) >= 8
    for name in single:
        for a, b in zip(single[name], paired[name]):
            assert torch.isfinite(a.float()).all(), name
            assert torch.equal(a, b), f"{name}: max |d| = {float((a.float() - b.float()).abs().max())}"
