import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


HP = dict(inter_channels=192, hidden_channels=192, filter_channels=768, n_heads=2, n_layers=6, kernel_size=3,
          p_dropout=0.1, resblock="1", resblock_kernel_sizes=[3, 7, 11],
          resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], upsample_rates=[8, 8, 2, 2],
          upsample_initial_channel=512, upsample_kernel_sizes=[16, 16, 4, 4])


@pytest.fixture(scope="session")
def hp():
    return dict(HP)


@pytest.fixture(scope="session")
def built_lib():
    """The in-tree shared library (built by __graft_entry__.build())."""
    from music_generation_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib.lib()
