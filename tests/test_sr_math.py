"""Stochastic rounding as utils/sr.py states it (the bit-exact oracle of csrc/sr.cuh): the
properties the optimizer relies on, checked on the CPU."""
import math

import torch

from collective_communication_mpi_b200.utils import sr


def _neighbours(x: torch.Tensor):
    """bf16 values just below / above (or equal to) each fp32 x, as fp32."""
    u = x.view(torch.int32)
    lo = (u & ~0xFFFF).view(torch.float32)                      # truncation: towards zero
    hi = ((u & ~0xFFFF) + 0x10000).view(torch.float32)          # one bf16 ulp further from zero (sign-magnitude)
    exact = (u & 0xFFFF) == 0
    return lo, torch.where(exact, lo, hi)


def test_same_index_step_seed_gives_the_same_bits_and_different_ones_differ():
    idx = torch.arange(0, 4096)
    a = sr.bits(idx, step=3, seed=17)
    assert torch.equal(a, sr.bits(idx, step=3, seed=17))
    assert a.min() >= 0 and a.max() < 2 ** 32
    for other in (sr.bits(idx, step=4, seed=17), sr.bits(idx, step=3, seed=18), sr.bits(idx + 4096, step=3, seed=17)):
        assert (a == other).float().mean() < 0.01
    # bits are spread: every one of the 32 positions is set about half the time
    for b in range(32):
        frac = ((a >> b) & 1).float().mean().item()
        assert 0.45 < frac < 0.55, (b, frac)


def test_indices_beyond_32_bits_do_not_alias():
    lo = sr.bits(torch.arange(0, 1024), 1, 0)
    hi = sr.bits(torch.arange(0, 1024) + (1 << 32), 1, 0)
    assert (lo == hi).float().mean() < 0.01


def test_result_is_one_of_the_two_neighbouring_bf16_values():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(1 << 14, generator=g) * torch.logspace(-6, 6, 1 << 14)
    idx = torch.arange(x.numel())
    y = sr.round_bf16(x, idx, step=1, seed=5).float()
    lo, hi = _neighbours(x)
    assert bool(((y == lo) | (y == hi)).all())
    assert bool((y == lo).any()) and bool((y == hi).any())


def test_values_already_in_bf16_are_unchanged_for_every_draw():
    x = torch.randn(4096).bfloat16().float()
    for step in range(5):
        assert torch.equal(sr.round_bf16(x, torch.arange(4096), step, 9).float(), x)


def test_non_finite_values_pass_through():
    x = torch.tensor([float("inf"), float("-inf"), float("nan"), 1.0, -0.0])
    y = sr.round_bf16(x, torch.arange(5), 2, 2).float()
    assert y[0] == float("inf") and y[1] == float("-inf") and math.isnan(y[2].item())
    assert y[3] == 1.0 and y[4] == 0.0 and math.copysign(1.0, y[4].item()) == -1.0


def test_rounding_is_unbiased_where_nearest_rounding_loses_the_update():
    """w = 1.0 has a bf16 ulp of 2^-7; an update of 2^-11 is lost by round-to-nearest every step.
    Stochastically the step down (2^-8, the ulp below 1.0) is taken with probability 1/8, so the MEAN moves by
    the true amount."""
    n = 1 << 16
    w = torch.ones(n)
    upd = 2.0 ** -11
    y = sr.round_bf16(w - upd, torch.arange(n), step=7, seed=1).float()
    assert torch.equal((w - upd).bfloat16().float(), w)                 # nearest: nothing happens
    moved = (y != 1.0).float().mean().item()
    # below 1.0 the ulp is 2^-8: the step down is taken with probability upd / 2^-8 = 1/8
    assert abs(moved - 0.125) < 0.01, moved
    assert abs(y.mean().item() - (1.0 - upd)) < 2e-5


def test_sgd_update_accumulates_small_steps_in_expectation():
    n, steps, lr = 1 << 14, 64, 1e-3
    w = torch.full((n,), 0.5).bfloat16()
    g = torch.full((n,), 0.25).bfloat16()                                 # step 2.5e-4; ulp(0.5) = 2^-8 = 3.9e-3
    for s in range(steps):
        w = sr.sgd_update(w, g, lr, base_index=0, step=s, seed=3)
    want = 0.5 - steps * lr * 0.25
    assert abs(w.float().mean().item() - want) < 3e-4
    nearest = torch.full((n,), 0.5).bfloat16()
    for s in range(steps):
        nearest = (nearest.float() - lr * g.float()).bfloat16()
    assert torch.equal(nearest.float(), torch.full((n,), 0.5))           # the failure SR is there to prevent


def test_base_index_shifts_the_draws_so_shards_of_one_tensor_are_consistent():
    """A weight updated as one run or as two runs with the right base indices gets the same bits:
    what lets the fused data-parallel update and the per-layer kernels agree."""
    g0 = torch.Generator().manual_seed(4)
    w = torch.randn(1000, generator=g0).bfloat16()
    g = torch.randn(1000, generator=g0).bfloat16()
    whole = sr.sgd_update(w, g, 1e-2, base_index=64, step=2, seed=8)
    a = sr.sgd_update(w[:400], g[:400], 1e-2, base_index=64, step=2, seed=8)
    b = sr.sgd_update(w[400:], g[400:], 1e-2, base_index=464, step=2, seed=8)
    assert torch.equal(whole, torch.cat([a, b]))
