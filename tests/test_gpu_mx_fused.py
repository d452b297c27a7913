"""Block-scaled fp8 (MXFP8) on the generic CTA-pair kernel: the CUDA quantiser against the torch statement of
the format, the plain GEMM against the product of the dequantised operands, and the two fused collectives
(all-gather -> GEMM with scale chunks pushed along, GEMM -> reduce-scatter) with emulated ranks on one GPU."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF = torch.bfloat16


def _rand(*shape, seed, scale=0.5):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).to(BF)


def _unswizzle(sf_chunked, R, K):
    """chunked hardware layout -> row-major [R, K/32] (inverse of ops.gemm_fp8.swizzle_scales)."""
    Rp, Kc = -(-R // 128) * 128, -(-K // 128)
    t = sf_chunked.reshape(Rp // 128, Kc, 32, 4, 4).permute(0, 3, 2, 1, 4).reshape(Rp, Kc * 4)
    return t[:R, :K // 32]


@pytest.mark.parametrize("R,K", [(128, 128), (256, 512), (200, 1024), (1024, 4096)])
def test_cuda_quantiser_matches_the_torch_statement_of_mxfp8(R, K):
    from collective_communication_mpi_b200.ops import gemm_fp8 as Q
    x = _rand(R, K, seed=R + K) * torch.logspace(-2, 2, R, device="cuda").unsqueeze(1).to(BF)
    q, sf = Q.quantize_mxfp8_cuda(x)
    q_ref, sf_ref = Q.quantize_mxfp8(x.float())
    assert torch.equal(_unswizzle(sf, R, K), sf_ref), "scales"
    assert torch.equal(q.view(torch.float8_e4m3fn).float(), q_ref.float()), "e4m3 values"
    assert torch.equal(sf, Q.swizzle_scales(sf_ref)) or R % 128 != 0        # (ragged rows: padding stays at the fill value)


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (256, 256, 512), (512, 384, 1024), (384, 128, 256), (1024, 2048, 2048)])
def test_block_scaled_gemm_on_the_generic_kernel(M, N, K):
    from collective_communication_mpi_b200.ops import gemm_fp8 as Q
    a = _rand(M, K, seed=1) * torch.logspace(-2, 2, M, device="cuda").unsqueeze(1).to(BF)
    b = _rand(N, K, seed=2)
    aq, asf = Q.quantize_mxfp8_cuda(a)
    bq, bsf = Q.quantize_mxfp8_cuda(b)
    out = Q.gemm_mx(aq, asf, bq, bsf, out_dtype=torch.float32)
    deq = lambda q, sf, R: Q.dequantize_mxfp8(q.view(torch.float8_e4m3fn), _unswizzle(sf, R, K))    # noqa: E731
    ref = deq(aq, asf, M) @ deq(bq, bsf, N).t()
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=1e-3 * ref.abs().max().item())
    # and the dedicated kernel of round 1 agrees bit for bit (same MMA order, same tile)
    old = Q.gemm_mxfp8(aq.view(torch.float8_e4m3fn), asf, bq.view(torch.float8_e4m3fn), bsf, out_dtype=torch.float32)
    assert torch.equal(out, old)


@pytest.fixture(params=[2, 4])
def emu(request):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr = request.param
    heap = SymmetricHeap.create_emulated(nr, 256 << 20)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    yield nr, heap, comms
    heap.close()


@pytest.mark.parametrize("M,N,K", [(256, 128, 1024), (640, 384, 2048), (200, 256, 1024)])
def test_allgather_gemm_block_scaled(emu, M, N, K):
    from collective_communication_mpi_b200.ops import gemm_fp8 as Q
    nr, heap, comms = emu
    chunk = K // nr
    U8 = torch.uint8
    xq_off, sfl_off = heap.alloc(M * chunk), heap.alloc(Q.sf_bytes(M, chunk))
    xf = [heap.alloc(M * K) for _ in range(2)]
    sff = [heap.alloc(Q.sf_bytes(M, K)) for _ in range(2)]
    ws = [Q.QuantizedWeight(_rand(K, N, seed=20 + r, scale=0.2)) for r in range(nr)]
    deq_w = [Q.dequantize_mxfp8(w.q.view(torch.float8_e4m3fn), _unswizzle(w.sf, N, K)) for w in ws]       # [N, K]
    for epoch in (1, 2, 3):
        shards = []
        for r in range(nr):
            x = _rand(M, chunk, seed=10 + r) * float(epoch) * torch.logspace(-1, 1, M, device="cuda").unsqueeze(1).to(BF)
            heap.view(sfl_off, (Q.sf_bytes(M, chunk),), U8, rank=r).fill_(127)
            q, sf = Q.quantize_mxfp8_cuda(x, heap.view(xq_off, (M, chunk), U8, rank=r), heap.view(sfl_off, (Q.sf_bytes(M, chunk),), U8, rank=r))
            shards.append(Q.dequantize_mxfp8(q.view(torch.float8_e4m3fn), _unswizzle(sf, M, chunk)))
        x_full = torch.cat(shards, dim=1)
        outs = [torch.zeros(M, N, device="cuda", dtype=torch.float32) for _ in range(nr)]
        Q.ag_gemm_mx_raw(comms, xq_off, chunk, sfl_off, xf[epoch & 1], sff[epoch & 1], [w.q for w in ws], [w.sf for w in ws], outs,
                         M, N, K, flag_base=0, epoch=epoch)
        torch.cuda.synchronize()
        for r in range(nr):
            ref = x_full @ deq_w[r].t()
            torch.testing.assert_close(outs[r], ref, rtol=2e-3, atol=1e-3 * ref.abs().max().item())


@pytest.mark.parametrize("M,chunk,K", [(256, 128, 512), (512, 256, 1024)])
def test_gemm_reduce_scatter_block_scaled(emu, M, chunk, K):
    from collective_communication_mpi_b200.ops import gemm_fp8 as Q
    nr, heap, comms = emu
    N = chunk * nr
    slots = [heap.alloc(nr * M * chunk * 2) for _ in range(2)]
    As = [Q.quantize_mxfp8_cuda(_rand(M, K, seed=30 + r)) for r in range(nr)]
    Ws = [Q.QuantizedWeight(_rand(K, N, seed=40 + r, scale=0.2)) for r in range(nr)]
    parts = []
    for (aq, asf), w in zip(As, Ws):
        a = Q.dequantize_mxfp8(aq.view(torch.float8_e4m3fn), _unswizzle(asf, M, K))
        b = Q.dequantize_mxfp8(w.q.view(torch.float8_e4m3fn), _unswizzle(w.sf, N, K))
        parts.append((a @ b.t()).to(BF).float())                   # each partial rounded to the bf16 slot type
    total = sum(parts)
    scale = torch.stack([p_.abs() for p_ in parts]).amax(0)
    for epoch in (1, 2):
        res = [_rand(M, chunk, seed=70 + r) for r in range(nr)] if epoch == 2 else None
        outs = [torch.zeros(M, chunk, device="cuda", dtype=BF) for _ in range(nr)]
        Q.gemm_rs_mx_raw(comms, [a for a, _ in As], [s for _, s in As], [w.q for w in Ws], [w.sf for w in Ws], slots[epoch & 1], outs,
                         M, N, K, flag_base=4096, epoch=epoch, residual=res)
        torch.cuda.synchronize()
        for r in range(nr):
            sl = slice(r * chunk, (r + 1) * chunk)
            ref = total[:, sl] + (res[r].float() if res is not None else 0.0)
            big = torch.maximum(scale[:, sl], ref.abs()).clamp_min(2.0 ** -20)
            ulp = torch.exp2(torch.floor(torch.log2(big)) - 7)
            err = (outs[r].float() - ref.to(BF).float()).abs()
            assert bool((err <= (nr + 1) * ulp).all()), f"rank {r}: {(err / ulp).max().item():.1f} ulps off"
