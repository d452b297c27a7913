"""Closing fused ops on real GPUs: flag words and symmetric buffers come back and the next owner
of a flag region starts clean (tests/workers/w_slot_reuse.py). Needs >= 2 GPUs. The CPU half of
the same feature is tests/test_flag_allocator.py."""
import pytest

from conftest import gpu_count, run_ranks

pytestmark = pytest.mark.gpu


def test_flag_regions_are_reusable_after_close():
    n = gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    ranks = 4 if n >= 4 else 2
    proc = run_ranks(ranks, ["tests/workers/w_slot_reuse.py"], transport="device")
    assert proc.returncode == 0, f"rc={proc.returncode}\n--- stdout\n{proc.stdout[-4000:]}\n--- stderr\n{proc.stderr[-3000:]}"
    assert f"SLOT REUSE OK ranks={ranks}" in proc.stdout
