"""Weight-gradient GEMM with the SGD step in its epilogue (csrc/gemm_sgd.cu) vs fp32 PyTorch.
First run on a B200 in round 2 (profiles/r2_unverified_1gpu.txt). The layer-level logic (dx before the
in-place weight update) is also covered on the CPU by tests/test_multirank_host.py."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF = torch.bfloat16


@pytest.mark.parametrize("T,M,N", [(256, 128, 128), (1024, 384, 640), (4096, 1024, 2048), (512, 200, 136)])
@pytest.mark.parametrize("layout", ["tn", "nn", "nt"])
def test_update_matches_fp32_reference(T, M, N, layout):
    from collective_communication_mpi_b200.ops.gemm import gemm_sgd_
    g = torch.Generator(device="cuda").manual_seed(0)
    w = (torch.randn(M, N, device="cuda", generator=g) * 0.1).to(BF)
    x = (torch.randn(T, M, device="cuda", generator=g) * 0.5).to(BF)
    dy = (torch.randn(T, N, device="cuda", generator=g) * 0.5).to(BF)
    lr = 1e-2
    ref = w.float() - lr * (x.float().t() @ dy.float())
    w0 = w.clone()
    if layout == "tn":        # x^T dy with both operands as stored by the layers
        gemm_sgd_(w, x, dy, lr, a_km=True, b_kn=True)
    elif layout == "nn":
        gemm_sgd_(w, x.t().contiguous(), dy, lr, b_kn=True)
    else:
        gemm_sgd_(w, x.t().contiguous(), dy.t().contiguous(), lr)
    torch.cuda.synchronize()
    assert not torch.equal(w, w0)
    # one bf16 rounding of the result: half an ulp of |w| plus the GEMM's own accumulation noise
    tol = ref.abs().max().item() * 2 ** -8 + 1e-3
    assert (w.float() - ref).abs().max().item() < tol


def test_train_step_with_fused_sgd_follows_the_unfused_one(monkeypatch):
    from mpi4py import MPI
    from collective_communication_mpi_b200.models import TrainStep, TransformerFCBlock
    comm = MPI.COMM_WORLD
    if comm.Get_size() != 1:
        pytest.skip("single-process test")
    T, H = 512, 256
    g = torch.Generator(device="cuda").manual_seed(1)
    x = (torch.randn(T, H, device="cuda", generator=g) * 0.5).to(BF)
    t = (torch.randn(T, H, device="cuda", generator=g) * 0.5).to(BF)
    losses = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("CCMPI_FUSED_SGD", flag)
        step = TrainStep(TransformerFCBlock(comm, T, H, 4, seed=1), comm, lr=1e-2)
        assert step.fused_sgd == (flag == "1")
        losses[flag] = [float(step(x, t)) for _ in range(5)]
    for a, b in zip(losses["0"], losses["1"]):
        assert abs(a - b) < 0.02 * max(abs(a), 1e-3), (losses["0"], losses["1"])
    assert losses["1"][-1] < losses["1"][0]
