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


# ---- stochastic rounding (csrc/sr.cuh): the kernels against the same arithmetic in torch (utils/sr.py) ----
def test_standalone_update_kernel_matches_the_torch_statement_bit_for_bit():
    from collective_communication_mpi_b200.ops import elementwise as EW
    from collective_communication_mpi_b200.utils import sr
    g = torch.Generator(device="cuda").manual_seed(3)
    for n, base, step, seed in ((1 << 16, 0, 0, 1), (8 * 1000, 123456, 7, 99), (1 << 20, (1 << 33) + 64, 2 ** 31 + 5, 0xDEADBEEF)):
        w = (torch.randn(n, device="cuda", generator=g) * 0.02).to(BF)
        gr = (torch.randn(n, device="cuda", generator=g) * 0.01).to(BF)
        want = sr.sgd_update(w, gr, 1e-3, base, step & 0xFFFFFFFF, seed)
        got = EW.sgd_sr_(w.clone(), gr, 1e-3, base_index=base, step=step, seed=seed)
        assert torch.equal(got, want), f"n={n}: {(got != want).sum().item()} of {n} elements differ"
        assert 0.02 < (got != w).float().mean().item() < 0.9, "a sub-ulp step must move some weights, not all"


def test_stochastic_rounding_is_unbiased_where_round_to_nearest_loses_the_step():
    from collective_communication_mpi_b200.ops import elementwise as EW
    g = torch.Generator(device="cuda").manual_seed(4)
    n = 1 << 18
    w0 = (torch.randn(n, device="cuda", generator=g) * 0.02).to(BF)
    gr = torch.full((n,), 0.01, device="cuda", dtype=BF)
    lr, steps = 1e-3, 400                        # each step is ~1e-5: far below half an ulp of most weights
    w = w0.clone()
    for s in range(steps):
        EW.sgd_sr_(w, gr, lr, step=s, seed=11)
    moved = (w.float() - w0.float()).mean().item()
    want = -lr * 0.01 * steps
    assert abs(moved - want) < 0.05 * abs(want), (moved, want)
    rn = w0.clone()
    for s in range(steps):
        rn = (rn.float() - lr * gr.float()).to(BF)
    assert abs((rn.float() - w0.float()).mean().item()) < 0.35 * abs(want), "round-to-nearest should lose most of it"


def test_gemm_sgd_epilogue_with_stochastic_rounding():
    from collective_communication_mpi_b200.ops.gemm import gemm, gemm_sgd_
    from collective_communication_mpi_b200.utils import sr
    g = torch.Generator(device="cuda").manual_seed(5)
    M, Nn, K = 512, 384, 256
    x = (torch.randn(K, M, device="cuda", generator=g) * 0.5).to(BF)      # [tokens, in]  -> a_km
    dy = (torch.randn(K, Nn, device="cuda", generator=g) * 0.5).to(BF)    # [tokens, out] -> b_kn
    w0 = (torch.randn(M, Nn, device="cuda", generator=g) * 0.02).to(BF)
    acc = gemm(x, dy, a_km=True, b_kn=True, out_dtype=torch.float32)      # the kernel's own fp32 accumulator
    idx = torch.arange(M, device="cuda", dtype=torch.int64).view(-1, 1) * Nn + torch.arange(Nn, device="cuda", dtype=torch.int64)
    lr32 = torch.tensor(1e-4, dtype=torch.float32, device="cuda")
    want = sr.round_bf16(w0.float() - lr32 * acc, idx, 9, 21)
    got = gemm_sgd_(w0.clone(), x, dy, 1e-4, a_km=True, b_kn=True, step=9, seed=21)
    assert torch.equal(got, want), f"{(got != want).sum().item()} of {got.numel()} elements differ"
    rn = gemm_sgd_(w0.clone(), x, dy, 1e-4, a_km=True, b_kn=True)         # step=None: round to nearest
    assert torch.equal(rn, (w0.float() - lr32 * acc).to(BF))
