"""All-gather->GEMM launches on ONE flag region whose CTA count changes from call to call (another
N, another SM budget): the ready flags count pushing CTAs, so the host passes the running sum of
CTA counts as an absolute target (ops.fused.ag_ctas; ADVICE round 1, ops/fused.py). Runs in its own
process with a short device watchdog: a protocol error shows up as a trap there, not as a dead CUDA
context in the test session."""
import os
import subprocess
import sys

import pytest

from conftest import REPO, gpu_count

pytestmark = pytest.mark.gpu


def test_ready_flag_targets_follow_a_changing_cta_count():
    if gpu_count() < 1:
        pytest.skip("needs a GPU")
    env = dict(os.environ, PYTHONPATH=REPO, CCMPI_DEVICE_TIMEOUT_S="10")
    p = subprocess.run([sys.executable, "tests/workers/w_flag_targets.py"], cwd=REPO, env=env, capture_output=True,
                       text=True, timeout=300)
    assert p.returncode == 0, f"rc={p.returncode}\n--- stdout\n{p.stdout[-3000:]}\n--- stderr\n{p.stderr[-3000:]}"
    assert "FLAG TARGETS OK" in p.stdout
