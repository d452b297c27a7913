"""Host-side halves of the fixes for round 1's advisor findings, checked without a GPU.
Device halves: tests/test_zzz_flag_targets.py, tests/test_gpu_multirank.py."""
import pytest

from conftest import run_ranks


def _ok(proc, token):
    assert proc.returncode == 0, f"rc={proc.returncode}\n--- stdout\n{proc.stdout[-4000:]}\n--- stderr\n{proc.stderr[-3000:]}"
    assert token in proc.stdout, proc.stdout[-2000:]


@pytest.mark.parametrize("n", [2, 4])
def test_zero_copy_route_needs_the_same_heap_offset_on_every_member(n):
    _ok(run_ranks(n, ["tests/workers/w_route_agreement.py"], timeout=120), f"ROUTE AGREEMENT OK ranks={n}")


def test_ready_flag_contributions_per_launch_depend_on_shape_and_sm_budget():
    """What an all-gather->GEMM launch adds to each ready flag is its CTA count; it changes with N and
    with the SM budget, which is why the ops keep a running sum (absolute target) instead of
    ``epoch * ctas``."""
    from collective_communication_mpi_b200.ops.fused import ag_ctas
    M = 8192
    full = ag_ctas(0, M, 8192, 1, 148)
    assert full == 148                                      # more tiles than SMs: every SM gets a CTA
    assert ag_ctas(0, M, 8192, 1, 132) == 132               # backward kernels sized to the SMs a DP bucket leaves
    assert ag_ctas(0, M, 8192, 2, 148) == 74                # two emulated ranks share the GPU
    small = ag_ctas(0, 512, 256, 1, 148)
    assert 0 < small < 148                                  # fewer tiles than SMs
    assert ag_ctas(0, 512, 256, 1, 148) == small            # deterministic: every member computes the same number
    seq = [ag_ctas(0, 512, N, 2, sms) for N, sms in [(256, 148), (1024, 148), (128, 148), (1024, 32)]]
    running = [sum(seq[:i + 1]) for i in range(len(seq))]
    assert running != [(i + 1) * seq[i] for i in range(len(seq))]   # the old target would have been wrong here
    with pytest.raises(Exception, match="unknown kernel kind|bad arguments"):
        ag_ctas(7, 512, 256, 1, 148)


def test_device_watchdog_setter_validates_before_touching_a_device():
    from collective_communication_mpi_b200 import _native as N
    lib = N.device_lib()
    assert lib.ccmpi_set_spin_timeout_ns(10) < 0            # below 1 ms: refused
    assert b"below 1 ms" in lib.ccmpi_last_error()


def test_device_watchdog_budget_comes_from_the_environment(monkeypatch):
    import contextlib
    import torch
    from collective_communication_mpi_b200 import _native as N
    calls = []

    class Lib:
        @staticmethod
        def ccmpi_set_spin_timeout_ns(ns):
            calls.append(ns)
            return 0

    monkeypatch.setattr(N, "device_lib", lambda: Lib)
    monkeypatch.setattr(torch.cuda, "device", lambda d: contextlib.nullcontext())
    monkeypatch.setattr(N, "_timeout_done", set())
    assert N.apply_device_timeout(0) == 120.0               # default matches the host plane's CCMPI_TIMEOUT
    assert N.apply_device_timeout(0) == -1.0                # once per device and process
    monkeypatch.setenv("CCMPI_DEVICE_TIMEOUT_S", "7.5")
    assert N.apply_device_timeout(1) == 7.5
    assert N.apply_device_timeout(1, seconds=2) == 2        # explicit value always applies
    assert calls == [120_000_000_000, 7_500_000_000, 2_000_000_000]
