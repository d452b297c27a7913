"""GEMMs with elementwise work fused into their epilogue (EPI_GELU_DUAL / EPI_MUL_DGELU / EPI_ADD in
csrc/gemm_kernel.cuh) against the unfused math in fp32 PyTorch: plain GEMM, all-gather -> GEMM and
GEMM -> reduce-scatter (emulated ranks on one GPU), and the kernel trace they can record."""
import math

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF = torch.bfloat16
C_, A_ = math.sqrt(2.0 / math.pi), 0.044715


def _rand(*shape, seed, scale=0.5):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).to(BF)


def _gelu(x):
    return 0.5 * x * (1.0 + torch.tanh(C_ * (x + A_ * x ** 3)))


def _dgelu(x):
    th = torch.tanh(C_ * (x + A_ * x ** 3))
    return 0.5 * (1.0 + th) + 0.5 * x * (1.0 - th * th) * C_ * (1.0 + 3.0 * A_ * x * x)


def _close(got, ref, what):
    # one bf16 rounding of the result (2^-8 relative) plus the SFU tanh (2^-11); absolute floor for values near zero
    err = (got.float() - ref).abs()
    bound = 6e-3 * ref.abs() + 4e-3 * max(1.0, ref.abs().max().item() / 8)
    assert bool((err <= bound).all()), f"{what}: max err {err.max().item():.4g} (|ref| max {ref.abs().max().item():.3g})"


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (384, 320, 512), (1024, 2048, 512), (136, 96, 1024)])
def test_plain_gemm_epilogues(M, N, K):
    from collective_communication_mpi_b200.ops.gemm import gemm
    x, w = _rand(M, K, seed=1), _rand(K, N, seed=2, scale=0.2)
    ref = x.float() @ w.float()
    u, act = gemm(x, w, b_kn=True, gelu_out=True)
    _close(u, ref, "pre-activation")
    _close(act, _gelu(u.float()), "gelu of the rounded pre-activation")
    du = gemm(x, w, b_kn=True, dgelu_of=u)
    _close(du, ref * _dgelu(u.float()), "product times gelu'(u)")
    r = _rand(M, N, seed=3)
    _close(gemm(x, w, b_kn=True, add=r), ref + r.float(), "product plus residual")
    # strided second operand (a column block of a wider tensor)
    wide = _rand(M, 2 * N, seed=4)
    _close(gemm(x, w, b_kn=True, add=wide[:, N:]), ref + wide[:, N:].float(), "strided residual")


def test_epilogue_argument_checks():
    from collective_communication_mpi_b200 import _native as Nat
    from collective_communication_mpi_b200.ops.gemm import gemm
    x, w = _rand(128, 64, seed=1), _rand(64, 128, seed=2)
    with pytest.raises(ValueError, match="at most one"):
        gemm(x, w, b_kn=True, gelu_out=True, add=_rand(128, 128, seed=3))
    with pytest.raises(ValueError, match="bfloat16"):
        gemm(x, w, b_kn=True, add=torch.zeros(128, 128, device="cuda"))
    with pytest.raises(Nat.NativeError, match="bf16 output"):
        gemm(x, w, b_kn=True, out_dtype=torch.float32, add=_rand(128, 128, seed=3))


@pytest.fixture(params=[2, 4])
def emu(request):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr = request.param
    heap = SymmetricHeap.create_emulated(nr, 256 << 20)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    yield nr, heap, comms
    heap.close()


@pytest.mark.parametrize("M,N,K,b_kn", [(256, 256, 512, True), (640, 384, 1024, False)])
def test_allgather_gemm_with_epilogues(emu, M, N, K, b_kn):
    from collective_communication_mpi_b200.ops import fused as F
    from collective_communication_mpi_b200.ops.gemm import EPI_ADD, EPI_GELU_DUAL, EPI_MUL_DGELU
    nr, heap, comms = emu
    chunk = K // nr
    x_off = heap.alloc(M * chunk * 2)
    xf = [heap.alloc(M * K * 2) for _ in range(2)]
    ws = [_rand(K, N, seed=20 + r, scale=0.2) for r in range(nr)]
    B = [w if b_kn else w.t().contiguous() for w in ws]
    for r in range(nr):
        heap.view(x_off, (M, chunk), BF, rank=r).copy_(_rand(M, chunk, seed=10 + r))
    x_full = torch.cat([heap.view(x_off, (M, chunk), BF, rank=r) for r in range(nr)], dim=1).float()
    refs = [x_full @ w.float() for w in ws]
    epoch = 0

    def run(op, e_in=None, e_out2=None):
        nonlocal epoch
        epoch += 1
        outs = [torch.zeros(M, N, device="cuda", dtype=BF) for _ in range(nr)]
        F.ag_gemm_raw(comms, x_off, chunk, xf[epoch & 1], B, outs, M, N, K, b_kn=b_kn, flag_base=0, epoch=epoch,
                      use_mc=False, epi_op=op, epi_in=e_in, epi_out2=e_out2)
        torch.cuda.synchronize()
        return outs

    acts = [torch.zeros(M, N, device="cuda", dtype=BF) for _ in range(nr)]
    us = run(EPI_GELU_DUAL, e_out2=acts)
    for r in range(nr):
        _close(us[r], refs[r], "u")
        _close(acts[r], _gelu(us[r].float()), "gelu(u)")
    dus = run(EPI_MUL_DGELU, e_in=us)
    res = [_rand(M, N, seed=90 + r) for r in range(nr)]
    sums = run(EPI_ADD, e_in=res)
    for r in range(nr):
        _close(dus[r], refs[r] * _dgelu(us[r].float()), "times gelu'")
        _close(sums[r], refs[r] + res[r].float(), "plus residual")


@pytest.mark.parametrize("M,chunk,K,b_kn", [(256, 256, 512, False), (1024, 128, 1024, True)])
def test_gemm_reduce_scatter_with_residual_and_rounded_partials_oracle(emu, M, chunk, K, b_kn):
    """The kernel rounds every rank's partial tile to bf16 once (the slot), sums the slots in fp32 in rank
    order, adds the residual and rounds once more. The oracle does exactly that, so the only legitimate
    differences are rounding flips of a partial (its fp32 value depends on the accumulation order inside
    the GEMM): at most one bf16 ulp per partial, plus one for the result. A missing or doubled partial
    would be ~128 ulps."""
    from collective_communication_mpi_b200.ops import fused as F
    nr, heap, comms = emu
    N = chunk * nr
    slots = [heap.alloc(nr * M * chunk * 2) for _ in range(2)]
    As = [_rand(M, K, seed=30 + r) for r in range(nr)]
    Ws = [_rand(K, N, seed=40 + r, scale=0.2) for r in range(nr)]
    B = [w if b_kn else w.t().contiguous() for w in Ws]
    partials = [(a.float() @ w.float()).to(BF).float() for a, w in zip(As, Ws)]     # rounded partials, rank order
    total = torch.zeros_like(partials[0])
    for p_ in partials:
        total += p_
    scale = torch.stack([p_.abs() for p_ in partials]).amax(0)
    for epoch, with_res in ((1, False), (2, True)):
        res = [_rand(M, chunk, seed=70 + r) for r in range(nr)] if with_res else None
        outs = [torch.zeros(M, chunk, device="cuda", dtype=BF) for _ in range(nr)]
        F.gemm_rs_raw(comms, As, B, slots[epoch & 1], outs, M, N, K, b_kn=b_kn, flag_base=4096, epoch=epoch, residual=res)
        torch.cuda.synchronize()
        for r in range(nr):
            sl = slice(r * chunk, (r + 1) * chunk)
            ref = total[:, sl] + (res[r].float() if with_res else 0.0)
            big = torch.maximum(scale[:, sl], ref.abs()).clamp_min(2.0 ** -20)
            ulp = torch.exp2(torch.floor(torch.log2(big)) - 7)
            err = (outs[r].float() - ref.to(BF).float()).abs()
            assert bool((err <= (nr + 1) * ulp).all()), f"rank {r}: {(err / ulp).max().item():.1f} ulps off"


def test_kernel_trace_records_every_role():
    from collective_communication_mpi_b200.ops.gemm import gemm
    from collective_communication_mpi_b200.utils import ktrace as KT
    x, w = _rand(4096, 1024, seed=1), _rand(1024, 4096, seed=2)
    buf = KT.new_buffer()
    gemm(x, w, b_kn=True, trace=buf)
    s = KT.summary(buf)
    assert s["ctas"] == 148 or s["ctas"] % 2 == 0 and s["ctas"] > 0
    assert s["kernel_us"] > 0 and s["tiles_per_cta"] >= 1
    for k in ("prod_empty_us", "mma_full_us", "epi_acc_us"):
        assert k in s and s[k]["max"] >= 0
