"""Fused elementwise kernels (csrc/elementwise.cu) vs their fp32 PyTorch statements."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF = torch.bfloat16


def _rand(*shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).to(BF)


@pytest.mark.parametrize("T,H", [(8, 8), (300, 256), (4096, 2048)])
def test_gate_forward_and_backward(T, H):
    from collective_communication_mpi_b200.ops import elementwise as EW
    qkv = _rand(T, 3 * H, seed=1)
    da = _rand(T, H, seed=2)
    a = EW.gate_forward(qkv)
    torch.testing.assert_close(a.float(), EW.gate_reference(qkv.float()), rtol=1e-2, atol=1e-2)
    d = EW.gate_backward(da, qkv)
    torch.testing.assert_close(d.float(), EW.gate_backward_reference(da.float(), qkv.float()), rtol=1e-2, atol=1e-2)
    # strided input: a column slice of a wider buffer
    wide = _rand(T, 3 * H + 64, seed=3)
    view = wide[:, 32:32 + 3 * H]                         # 64-byte offset, row stride 3H + 64
    torch.testing.assert_close(EW.gate_forward(view).float(), EW.gate_reference(view.float()), rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("shape", [(16,), (257, 64), (4096, 8192)])
def test_gelu_forward_and_backward(shape):
    from collective_communication_mpi_b200.ops import elementwise as EW
    u = _rand(*shape, seed=4, scale=3.0)
    dy = _rand(*shape, seed=5)
    torch.testing.assert_close(EW.gelu_forward(u).float(), EW.gelu_reference(u.float()), rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(EW.gelu_backward(dy, u).float(), EW.gelu_backward_reference(dy.float(), u.float()),
                               rtol=1e-2, atol=1e-2)
    big = torch.tensor([-30.0, -9.0, 9.0, 30.0, 0.0, 1e4, -1e4, 0.5], device="cuda", dtype=BF)
    assert torch.isfinite(EW.gelu_forward(big).float()).all() and torch.isfinite(EW.gelu_backward(big, big).float()).all()


@pytest.mark.parametrize("shape", [(8,), (1000, 256), (8192, 4096)])
def test_mse_loss_and_gradient_twice_in_a_row(shape):
    from collective_communication_mpi_b200.ops import elementwise as EW
    y, t = _rand(*shape, seed=6), _rand(*shape, seed=7)
    ref_loss, ref_dy = EW.mse_reference(y.float(), t.float())
    for _ in range(2):                                   # the workspace must be left ready for the next call
        loss, dy = EW.mse_loss_and_grad(y, t)
        torch.testing.assert_close(loss, ref_loss, rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(dy.float(), ref_dy, rtol=1e-2, atol=1e-9)
    again, _ = EW.mse_loss_and_grad(y, t)
    assert torch.equal(again, loss), "fixed-order reduction: same inputs, same bits"


def test_block_with_fused_elementwise_matches_eager_block(monkeypatch):
    """The whole training step with the switch on vs off: same loss trajectory within bf16 noise."""
    from mpi4py import MPI
    from collective_communication_mpi_b200.models import TrainStep, TransformerFCBlock
    comm = MPI.COMM_WORLD
    if comm.Get_size() != 1:
        pytest.skip("single-process test")
    T, H = 512, 256
    x, t = _rand(T, H, seed=8, scale=0.5), _rand(T, H, seed=9, scale=0.5)
    losses = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("CCMPI_FUSED_ELEMENTWISE", flag)
        step = TrainStep(TransformerFCBlock(comm, T, H, 4, seed=1), comm, lr=1e-2)
        losses[flag] = [float(step(x, t)) for _ in range(4)]
    for a, b in zip(losses["0"], losses["1"]):
        assert abs(a - b) < 0.02 * max(abs(a), 1e-3), (losses["0"], losses["1"])
    assert losses["1"][-1] < losses["1"][0]
