"""The math the fused elementwise kernels implement (csrc/elementwise.cu), checked on the CPU: the
hand-derived backward formulas in ops/elementwise.py must agree with autograd through the plain
PyTorch composition they replace, and the autograd wrappers must route gradients correctly."""
import pytest

torch = pytest.importorskip("torch")

from collective_communication_mpi_b200.ops import elementwise as EW  # noqa: E402


def test_gate_backward_formula_matches_autograd():
    torch.manual_seed(0)
    qkv = torch.randn(5, 7, 3 * 16, dtype=torch.float64, requires_grad=True)
    q, k, v = qkv.chunk(3, dim=-1)
    a = v * torch.sigmoid(q * k)
    da = torch.randn_like(a)
    a.backward(da)
    torch.testing.assert_close(EW.gate_reference(qkv.detach()), a.detach(), rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(EW.gate_backward_reference(da, qkv.detach()), qkv.grad, rtol=1e-5, atol=1e-6)


def test_gelu_formulas_match_torch_tanh_gelu():
    torch.manual_seed(1)
    u = (torch.randn(4, 33, dtype=torch.float64) * 3).requires_grad_(True)
    y = torch.nn.functional.gelu(u, approximate="tanh")
    dy = torch.randn_like(y)
    y.backward(dy)
    torch.testing.assert_close(EW.gelu_reference(u.detach()), y.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(EW.gelu_backward_reference(dy, u.detach()), u.grad, rtol=1e-5, atol=1e-6)


def test_mse_loss_and_grad_match_torch():
    torch.manual_seed(2)
    y = torch.randn(6, 40, dtype=torch.float64, requires_grad=True)
    t = torch.randn(6, 40, dtype=torch.float64)
    loss = torch.nn.functional.mse_loss(y, t)
    loss.backward()
    got_loss, got_dy = EW.mse_reference(y.detach(), t)
    torch.testing.assert_close(got_loss.double(), loss.detach(), rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(got_dy, y.grad, rtol=1e-5, atol=1e-7)


def test_autograd_wrappers_on_cpu_tensors():
    torch.manual_seed(3)
    qkv = torch.randn(3, 24, requires_grad=True)
    w = torch.randn(8, 8)
    out = EW.gelu(EW.gate(qkv) @ w).sum()
    out.backward()
    ref_in = qkv.detach().clone().requires_grad_(True)
    q, k, v = ref_in.chunk(3, dim=-1)
    ref = torch.nn.functional.gelu((v * torch.sigmoid(q * k)) @ w, approximate="tanh").sum()
    ref.backward()
    torch.testing.assert_close(out.detach(), ref.detach(), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(qkv.grad, ref_in.grad, rtol=1e-4, atol=1e-5)


def test_switch_defaults_to_on_and_zero_turns_it_off(monkeypatch):
    monkeypatch.delenv("CCMPI_FUSED_ELEMENTWISE", raising=False)
    assert EW.enabled() is True
    monkeypatch.setenv("CCMPI_FUSED_ELEMENTWISE", "0")
    assert EW.enabled() is False
