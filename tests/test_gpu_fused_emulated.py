"""Fused GEMM+collective kernels with emulated ranks on one GPU (one cooperative launch holds
every rank's CTAs). Oracle: the unfused math in fp32 PyTorch -- gather/concatenate, matmul,
rank-ordered sum, slice -- i.e. the reference's dataflow (func_impl.py:76-187) around GEMMs."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF = torch.bfloat16


@pytest.fixture(params=[2, 4])
def emu(request):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr = request.param
    heap = SymmetricHeap.create_emulated(nr, 256 << 20)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    yield nr, heap, comms
    heap.close()


def _rand(*shape, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * 0.5).to(BF)


@pytest.mark.parametrize("M,N,K,b_kn", [(256, 256, 512, True), (640, 384, 1024, True), (1024, 512, 2048, False),
                                        (200, 128, 512, True)])
def test_allgather_gemm_k_gather(emu, M, N, K, b_kn):
    from collective_communication_mpi_b200.ops import fused as F
    nr, heap, comms = emu
    chunk = K // nr
    x_off = heap.alloc(M * chunk * 2)
    xf = [heap.alloc(M * K * 2) for _ in range(2)]
    xs = [_rand(M, chunk, seed=10 + r) for r in range(nr)]
    ws = [_rand(K, N, seed=20 + r) for r in range(nr)]              # each rank has its own weight shard
    for epoch in (1, 2, 3):                                          # repeated launches: flags carry the epoch
        for r in range(nr):
            heap.view(x_off, (M, chunk), BF, rank=r).copy_(xs[r] * epoch)
        # the oracle sees exactly the bf16 values the kernel sees
        x_full = torch.cat([heap.view(x_off, (M, chunk), BF, rank=r) for r in range(nr)], dim=1).float()
        outs = [torch.zeros(M, N, device="cuda", dtype=torch.float32) for _ in range(nr)]
        B = [w if b_kn else w.t().contiguous() for w in ws]
        F.ag_gemm_raw(comms, x_off, chunk, xf[epoch & 1], B, outs, M, N, K, b_kn=b_kn, flag_base=0, epoch=epoch,
                      num_comm_ctas=4, use_mc=False)
        torch.cuda.synchronize()
        for r in range(nr):
            torch.testing.assert_close(outs[r], x_full @ ws[r].float(), rtol=1e-3, atol=2e-2)


@pytest.mark.parametrize("M,chunk,K,b_kn,dt", [(256, 256, 512, False, BF), (384, 128, 256, False, torch.float32),
                                               (1024, 256, 1024, True, BF), (200, 256, 512, False, BF)])
def test_gemm_reduce_scatter_n_scatter(emu, M, chunk, K, b_kn, dt):
    from collective_communication_mpi_b200.ops import fused as F
    nr, heap, comms = emu
    N = chunk * nr
    es = 2 if dt == BF else 4
    slots = [heap.alloc(nr * M * chunk * es) for _ in range(2)]
    As = [_rand(M, K, seed=30 + r) for r in range(nr)]
    Ws = [_rand(K, N, seed=40 + r) for r in range(nr)]
    # oracle: every rank's partial rounded to the slot dtype, summed in fp32 in rank order (what the kernel does)
    partials = [(a.float() @ w.float()).to(dt).float() for a, w in zip(As, Ws)]
    total = torch.zeros_like(partials[0])
    for p_ in partials:
        total += p_
    scale = torch.stack([p_.abs() for p_ in partials]).amax(0)
    for epoch in (1, 2):
        outs = [torch.zeros(M, chunk, device="cuda", dtype=dt) for _ in range(nr)]
        B = [w if b_kn else w.t().contiguous() for w in Ws]
        F.gemm_rs_raw(comms, As, B, slots[epoch & 1], outs, M, N, K, b_kn=b_kn, flag_base=4096, epoch=epoch)
        torch.cuda.synchronize()
        for r in range(nr):
            sl = slice(r * chunk, (r + 1) * chunk)
            ref = total[:, sl]
            if dt == BF:    # <= one bf16 ulp per partial (rounding flips) + one for the result; a lost partial is ~128
                big = torch.maximum(scale[:, sl], ref.abs()).clamp_min(2.0 ** -20)
                ulp = torch.exp2(torch.floor(torch.log2(big)) - 7)
                err = (outs[r].float() - ref.to(BF).float()).abs()
                assert bool((err <= (nr + 1) * ulp).all()), f"rank {r}: {(err / ulp).max().item():.1f} ulps off"
            else:
                torch.testing.assert_close(outs[r], ref, rtol=1e-4, atol=2e-3)
        # deterministic: same inputs, same bits
        again = [torch.zeros_like(o) for o in outs]
    F.gemm_rs_raw(comms, As, B, slots[1], again, M, N, K, b_kn=b_kn, flag_base=4096, epoch=3)
    torch.cuda.synchronize()
    for r in range(nr):
        assert torch.equal(again[r], outs[r])


@pytest.mark.parametrize("M,chunk,K", [(256, 128, 512), (512, 256, 1024)])
def test_gemm_allreduce(emu, M, chunk, K):
    from collective_communication_mpi_b200.ops import fused as F
    nr, heap, comms = emu
    N = chunk * nr
    slots = [heap.alloc(nr * M * chunk * 2) for _ in range(2)]
    out_off = [heap.alloc(M * N * 2) for _ in range(2)]
    As = [_rand(M, K, seed=50 + r) for r in range(nr)]
    Ws = [_rand(N, K, seed=60 + r) for r in range(nr)]
    partials = [(a.float() @ w.float().t()).to(BF).float() for a, w in zip(As, Ws)]     # bf16 slots, rank order
    total = torch.zeros_like(partials[0])
    for p_ in partials:
        total += p_
    scale = torch.stack([p_.abs() for p_ in partials]).amax(0)
    ulp = torch.exp2(torch.floor(torch.log2(torch.maximum(scale, total.abs()).clamp_min(2.0 ** -20))) - 7)
    for epoch in (1, 2):
        F.gemm_ar_raw(comms, As, Ws, slots[epoch & 1], out_off[epoch & 1], M, N, K, b_kn=False, flag_base=8192,
                      epoch=epoch, use_mc=False)
        torch.cuda.synchronize()
        got = [heap.view(out_off[epoch & 1], (M, N), BF, rank=r).float() for r in range(nr)]
        for r in range(nr):
            err = (got[r] - total.to(BF).float()).abs()
            assert bool((err <= (nr + 1) * ulp).all()), f"rank {r}: {(err / ulp).max().item():.1f} ulps off"
            assert torch.equal(got[r], got[0]), "all ranks hold the same bits"


@pytest.mark.parametrize("M,N,K,dt", [(256, 256, 512, BF), (384, 128, 256, torch.float32), (200, 200, 512, BF)])
def test_gemm_allgather_epilogue(emu, M, N, K, dt):
    from collective_communication_mpi_b200.ops import fused as F
    nr, heap, comms = emu
    es = 2 if dt == BF else 4
    y_off = heap.alloc(M * N * nr * es)
    As = [_rand(M, K, seed=70 + r) for r in range(nr)]
    Ws = [_rand(K, N, seed=80 + r) for r in range(nr)]
    ref = torch.cat([a.float() @ w.float() for a, w in zip(As, Ws)], dim=1)
    for _ in range(2):
        F.gemm_ag_raw(comms, As, Ws, y_off, M, N, K, b_kn=True, out_dtype=dt, use_mc=False)
    torch.cuda.synchronize()
    for r in range(nr):
        got = heap.view(y_off, (M, N * nr), dt, rank=r).float()
        tol = 0.15 if dt == BF else 2e-2
        assert (got - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item() / 10)


def test_shape_constraints_are_reported(emu):
    from collective_communication_mpi_b200 import _native as Nat
    from collective_communication_mpi_b200.ops import fused as F
    nr, heap, comms = emu
    a = _rand(128, 128, seed=1)
    outs = [torch.zeros(128, 64, device="cuda", dtype=BF) for _ in range(nr)]
    with pytest.raises(Nat.NativeError, match="multiple of 128"):
        F.gemm_rs_raw(comms, [a] * nr, [a] * nr, 0, outs, 128, 64 * nr, 128, b_kn=False, flag_base=0, epoch=1)
