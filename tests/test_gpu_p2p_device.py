"""NVLink route of Send/Recv for CUDA tensors (the default on the device plane). Needs >= 2 GPUs. The control
path is tested without GPUs in tests/test_p2p_device_control.py."""
import pytest

from conftest import gpu_count, run_ranks

pytestmark = pytest.mark.gpu


def test_cuda_tensors_travel_over_the_peer_mapping():
    n = gpu_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = min(n, 4)
    p = run_ranks(n, ["tests/workers/w_p2p_device.py"], transport="device", env={"CCMPI_P2P": "device"})
    assert p.returncode == 0, f"rc={p.returncode}\n{p.stdout[-3000:]}\n{p.stderr[-3000:]}"
    assert f"P2P DEVICE OK ranks={n}" in p.stdout
