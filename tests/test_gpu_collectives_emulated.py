"""Collective kernels with several *emulated* ranks on one GPU: all ranks' blocks run inside one
cooperative launch, so the cross-rank flag protocol is exercised without ever making two
kernels wait for each other on the same device."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(params=[2, 4])
def emu(request):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr = request.param
    heap = SymmetricHeap.create_emulated(nr, 64 << 20)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    yield nr, heap, comms
    heap.close()


def _fill(heap, off, nr, count, dt, seed=0):
    g = torch.Generator().manual_seed(seed)
    hosts = []
    for r in range(nr):
        x = (torch.randn(count, generator=g) * 4).to(dt) if dt.is_floating_point else \
            torch.randint(-1000, 1000, (count,), generator=g, dtype=dt)
        heap.view(off, (count,), dt, rank=r).copy_(x)
        hosts.append(x)
    return hosts


@pytest.mark.parametrize("dt,op,count,algo", [
    (torch.int64, "min", 100, "oneshot"), (torch.int64, "sum", 10, "oneshot"), (torch.int32, "max", 7, "auto"),
    (torch.float32, "sum", 1 << 16, "oneshot"), (torch.float32, "sum", 1 << 18, "twoshot"),
    (torch.bfloat16, "sum", 1 << 18, "twoshot"), (torch.float64, "max", 4096, "twoshot"),
    (torch.float16, "sum", 4096, "auto"), (torch.bfloat16, "sum", 1000, "auto"),
])
def test_allreduce_variants(emu, dt, op, count, algo):
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    es = torch.empty((), dtype=dt).element_size()
    src, dst, tmp = heap.alloc(count * es), heap.alloc(count * es), heap.alloc(count * es)
    hosts = _fill(heap, src, nr, count, dt)
    stack = torch.stack([h.double() if dt.is_floating_point else h for h in hosts])
    ref = {"sum": stack.sum(0), "min": stack.min(0).values, "max": stack.max(0).values}[op]
    for _ in range(3):  # repeated launches reuse the epoch-numbered flags
        CO.allreduce(comms, CO.cc_dtype(dt), {"sum": N.SUM, "min": N.MIN, "max": N.MAX}[op], src, dst, count,
                     tmp_off=tmp, algo={"auto": N.AR_AUTO, "oneshot": N.AR_ONESHOT, "twoshot": N.AR_TWOSHOT}[algo])
    torch.cuda.synchronize()
    outs = [heap.view(dst, (count,), dt, rank=r).cpu() for r in range(nr)]
    for o in outs[1:]:
        assert torch.equal(o, outs[0]), "every rank must hold bit-identical results"
    if dt.is_floating_point:
        tol = {torch.float32: 1e-4, torch.float64: 1e-9, torch.bfloat16: 0.25, torch.float16: 0.05}[dt]
        assert (outs[0].double() - ref).abs().max().item() <= tol
    else:
        assert torch.equal(outs[0], ref.to(dt))


def test_allreduce_scale_and_cast_for_dp_gradients(emu):
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    count = 1 << 16
    src, dst, tmp = heap.alloc(count * 4), heap.alloc(count * 2), heap.alloc(count * 4)
    hosts = _fill(heap, src, nr, count, torch.float32, seed=3)
    ref = torch.stack(hosts).sum(0) / nr
    for algo in (N.AR_ONESHOT, N.AR_TWOSHOT):
        CO.allreduce(comms, N.F32, N.SUM, src, dst, count, tmp_off=tmp, algo=algo, scale=1.0 / nr, out_dtype=N.BF16)
        torch.cuda.synchronize()
        for r in range(nr):
            got = heap.view(dst, (count,), torch.bfloat16, rank=r).cpu().float()
            assert (got - ref).abs().max().item() < 0.05


def test_allgather_flat_and_last_axis(emu):
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    for dt, per in ((torch.int64, 2), (torch.bfloat16, 12288), (torch.float32, 5)):
        es = torch.empty((), dtype=dt).element_size()
        s, d = heap.alloc(per * es), heap.alloc(per * es * nr)
        parts = []
        for r in range(nr):
            x = torch.arange(per, dtype=torch.float32).add(100 * r).to(dt)
            parts.append(x)
            heap.view(s, (per,), dt, rank=r).copy_(x)
        CO.allgather(comms, s, d, 1, per * es, 0, 0)
        torch.cuda.synchronize()
        for r in range(nr):
            assert torch.equal(heap.view(d, (per * nr,), dt, rank=r).cpu(), torch.cat(parts))
    b, s_, p = 4, 8, 2   # (b, s, part) float64 shards -> (b, s, part*nr)
    full = torch.arange(b * s_ * p * nr, dtype=torch.float64).reshape(b, s_, p * nr)
    so, do = heap.alloc(b * s_ * p * 8), heap.alloc(b * s_ * p * nr * 8)
    for r in range(nr):
        heap.view(so, (b, s_, p), torch.float64, rank=r).copy_(full[:, :, r * p:(r + 1) * p])
    CO.allgather(comms, so, do, b * s_, p * 8, p * 8, p * nr * 8)
    torch.cuda.synchronize()
    for r in range(nr):
        assert torch.equal(heap.view(do, (b, s_, p * nr), torch.float64, rank=r).cpu(), full)


def test_reduce_scatter_flat_and_last_axis(emu):
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    per = 2
    s, d = heap.alloc(per * nr * 8), heap.alloc(per * 8)
    srcs = _fill(heap, s, nr, per * nr, torch.int64, seed=5)
    CO.reduce_scatter(comms, N.I64, N.MIN, s, d, 1, per, 0, 0)
    torch.cuda.synchronize()
    ref = torch.stack(srcs).min(0).values
    for r in range(nr):
        assert torch.equal(heap.view(d, (per,), torch.int64, rank=r).cpu(), ref[r * per:(r + 1) * per])
    rows, w = 64, 256
    s, d = heap.alloc(rows * w * nr * 2), heap.alloc(rows * w * 2)
    g = torch.randn(nr, rows, w * nr).bfloat16()
    for r in range(nr):
        heap.view(s, (rows, w * nr), torch.bfloat16, rank=r).copy_(g[r])
    CO.reduce_scatter(comms, N.BF16, N.SUM, s, d, rows, w, w * nr * 2, w * 2)
    torch.cuda.synchronize()
    tot = g.float().sum(0)
    for r in range(nr):
        got = heap.view(d, (rows, w), torch.bfloat16, rank=r).cpu().float()
        assert (got - tot[:, r * w:(r + 1) * w]).abs().max().item() < 0.1


@pytest.mark.parametrize("seg,staggered", [(1, False), (1, True), (512, False), (512, True), (3, False)])
def test_alltoall_both_schedules(emu, seg, staggered):
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    s, d = heap.alloc(seg * nr * 8), heap.alloc(seg * nr * 8)
    for r in range(nr):
        heap.view(s, (seg * nr,), torch.int64, rank=r).copy_(torch.arange(seg * nr) + r * 100000)
    CO.alltoall(comms, s, d, seg * 8, staggered=staggered)
    CO.barrier(comms)
    torch.cuda.synchronize()
    for r in range(nr):
        got = heap.view(d, (nr, seg), torch.int64, rank=r).cpu()
        for j in range(nr):
            assert torch.equal(got[j], torch.arange(r * seg, (r + 1) * seg) + j * 100000)


def test_unsupported_requests_fail_loudly(emu):
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    off = heap.alloc(1024)
    with pytest.raises(N.NativeError, match="SUM, MIN, MAX"):
        CO.allreduce(comms, N.F32, 7, off, off, 16)
    with pytest.raises(N.NativeError, match="multicast"):
        CO.allreduce(comms, N.F32, N.SUM, off, off, 16, algo=N.AR_NVLS)
