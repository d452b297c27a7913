"""Real multi-GPU runs (one process per GPU, symmetric heap over NVLink). Needs >= 2 GPUs; on a
1-GPU box only the single-rank device path is exercised."""
import os

import pytest

from conftest import gpu_count, run_ranks

pytestmark = pytest.mark.gpu

# NumPy payloads below CCMPI_HOST_ROUTE_MAX would stay on the shared-memory plane (mpi/core.py:
# _DeviceSide.agree); these tests are about the kernels, so the route is off unless a test says otherwise
KERNELS = {"CCMPI_HOST_ROUTE_MAX": "0"}


def _ok(proc, token):
    assert proc.returncode == 0, f"rc={proc.returncode}\n--- stdout\n{proc.stdout[-4000:]}\n--- stderr\n{proc.stderr[-3000:]}"
    assert token in proc.stdout, proc.stdout[-2000:]


def _ranks_for_box(wanted):
    n = gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    ks = [k for k in wanted if k <= n]
    if os.environ.get("CCMPI_TEST_LARGEST_ONLY") == "1":   # budgeted 8-GPU runs: only the biggest group
        ks = ks[-1:]
    return ks


@pytest.mark.parametrize("cuda", [False, True])
def test_collectives_on_device_plane(cuda):
    for n in _ranks_for_box([2, 4, 8]):
        _ok(run_ranks(n, ["tests/workers/w_collectives.py"] + (["--cuda"] if cuda else []), transport="device", env=KERNELS),
            f"COLLECTIVES OK ranks={n} cuda={cuda}")


@pytest.mark.parametrize("cuda", [False, True])
def test_tp_steps_on_device_plane(cuda):
    for n in _ranks_for_box([4, 8]):
        _ok(run_ranks(n, ["tests/workers/w_tp_steps.py"] + (["--cuda"] if cuda else []), transport="device", env=KERNELS),
            f"TP STEPS OK ranks={n} cuda={cuda}")


def test_small_host_payloads_take_the_host_plane_by_default():
    """Default routing on a device-plane job: same checks, tiny NumPy payloads over shared memory,
    large ones and CUDA tensors through the kernels."""
    for n in _ranks_for_box([2, 4, 8]):
        _ok(run_ranks(n, ["tests/workers/w_collectives.py"], transport="device"), f"COLLECTIVES OK ranks={n} cuda=False")


def test_single_rank_device_world():
    if gpu_count() < 1:
        pytest.skip("needs a GPU")
    code = ("import torch\nfrom mpi4py import MPI\nc = MPI.COMM_WORLD\n"
            "x = torch.arange(10, device='cuda'); y = torch.empty_like(x)\n"
            "c.Allreduce(x, y); assert torch.equal(x, y); print('ONE OK')\n")
    _ok(run_ranks(1, ["-c", code], transport="device"), "ONE OK")


@pytest.mark.parametrize("mp,dp", [(2, 1), (1, 2), (2, 2), (4, 2), (2, 4), (8, 1)])
def test_tp_layers_and_train_step_match_dense_model(mp, dp):
    if gpu_count() < mp * dp:
        pytest.skip(f"needs {mp * dp} GPUs")
    _ok(run_ranks(mp * dp, ["tests/workers/w_layers.py", str(mp), str(dp)], transport="device"), f"LAYERS OK mp={mp} dp={dp}")


def test_single_gpu_train_step():
    if gpu_count() < 1:
        pytest.skip("needs a GPU")
    _ok(run_ranks(1, ["tests/workers/w_layers.py", "1", "1"], transport="device"), "LAYERS OK mp=1 dp=1")


def test_dp_allreduce_fused_with_sgd_update_on_every_group_size():
    """multimem.ld_reduce of the gradients + stochastic-rounding update + multimem.st of the weights."""
    for n in _ranks_for_box([2, 4, 8]):
        p = run_ranks(n, ["tests/workers/w_dp_update.py"], transport="device")
        assert p.returncode == 0, f"rc={p.returncode}\n{p.stdout[-3000:]}\n{p.stderr[-3000:]}"
        assert f"DP UPDATE OK ranks={n}" in p.stdout or "DP UPDATE SKIPPED" in p.stdout, p.stdout[-2000:]
