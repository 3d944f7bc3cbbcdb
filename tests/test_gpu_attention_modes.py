"""The attention entry point selects its kernel once per process (B200SVS_ATTN_TS, B200SVS_ATTN_TS_PARTS: DESIGN.md 4.2);
the default (attention_ts, Q / P operands in tensor memory, O accumulated in TMEM) is what every other GPU test runs.
Here the remaining selections run the same reference-semantics cases in a child process each: the round-1 kernel
(operands from shared memory), attention_ts for split launches only, the two-CTA-per-SM plain variant, four softmax
warps per lane quadrant."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("env", [{"B200SVS_ATTN_TS": "0"}, {"B200SVS_ATTN_TS": "1"}, {"B200SVS_ATTN_TS": "3"},
                                 {"B200SVS_ATTN_TS": "2", "B200SVS_ATTN_TS_PARTS": "4"}],
                         ids=["tc", "ts_split_only", "ts_two_ctas", "ts_four_parts"])
def test_attention_kernel_selections(env):
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_misc.py"), "-q", "-x", "-m", "gpu",
                        "-k", "attention_tc", "-p", "no:cacheprovider"], cwd=ROOT, env=dict(os.environ, **env),
                       capture_output=True, text=True, timeout=600)
    tail = (r.stdout + r.stderr)[-1500:]
    assert r.returncode == 0, tail
    assert " passed" in r.stdout and "failed" not in r.stdout, tail
