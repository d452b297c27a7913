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


@pytest.mark.parametrize("flags", [[], ["--sp"], ["--megatron"]], ids=["feature-sharded", "sequence-parallel", "megatron"])
def test_blocks_can_be_given_another_token_count_on_gpus(flags):
    """``set_tokens`` on the fused kernels: the old ops' flag words and symmetric buffers come back
    (the worker checks the allocators) and the rebuilt ops match the dense model."""
    if gpu_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    sp, meg = int("--sp" in flags), int("--megatron" in flags)
    proc = run_ranks(2, ["tests/workers/w_set_tokens.py", "2"] + flags, transport="device")
    assert proc.returncode == 0, f"rc={proc.returncode}\n--- stdout\n{proc.stdout[-4000:]}\n--- stderr\n{proc.stderr[-3000:]}"
    assert f"SET TOKENS OK mp=2 sp={sp} megatron={meg}" in proc.stdout


@pytest.mark.parametrize("mp,dp,depth", [(2, 1, 3), (2, 2, 2)])
def test_block_stack_shares_one_scatter_op_on_gpus(mp, dp, depth):
    """``TransformerFCStack`` on the fused kernels: 4 x depth layers launch their reduce-scatters through one op
    (one flag region, one pair of slot buffers); results match the dense model, a few steps train, close() of
    every holder releases the op once."""
    if gpu_count() < mp * dp:
        pytest.skip(f"needs {mp * dp} GPUs")
    proc = run_ranks(mp * dp, ["tests/workers/w_stack.py", str(mp), str(dp), str(depth)], transport="device")
    assert proc.returncode == 0, f"rc={proc.returncode}\n--- stdout\n{proc.stdout[-4000:]}\n--- stderr\n{proc.stderr[-3000:]}"
    assert f"STACK OK mp={mp} dp={dp} depth={depth}" in proc.stdout
