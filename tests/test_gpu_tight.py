"""Group-size-specialised reduce-scatter / two-shot all-reduce (csrc/collectives_tight.cu) against the regular
kernels, emulated ranks on one GPU: results must be bit-identical (same rank-ordered reduction)."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(params=[2, 4])
def emu(request):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr = request.param
    heap = SymmetricHeap.create_emulated(nr, 128 << 20)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    yield nr, heap, comms
    heap.close()


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("count", [4096, 1 << 20, (1 << 20) + 4096])
def test_two_shot_allreduce(emu, dt, count):
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    es = 4 if dt == torch.float32 else 2
    src, tmp = heap.alloc(count * es), heap.alloc(count * es)
    d0, d1 = heap.alloc(count * es), heap.alloc(count * es)
    g = torch.Generator(device="cuda").manual_seed(0)
    for r in range(nr):
        heap.view(src, (count,), dt, rank=r).copy_((torch.randn(count, device="cuda", generator=g) * (r + 1)).to(dt))
    cc = CO.cc_dtype(dt)
    for op in (N.SUM, N.MAX):
        CO.allreduce(comms, cc, op, src, d0, count, tmp_off=tmp, algo=N.AR_TWOSHOT, scale=0.5, tight=False)
        CO.allreduce(comms, cc, op, src, d1, count, tmp_off=tmp, algo=N.AR_TWOSHOT, scale=0.5, tight=True)
        torch.cuda.synchronize()
        for r in range(nr):
            assert torch.equal(heap.view(d0, (count,), dt, rank=r), heap.view(d1, (count,), dt, rank=r)), (op, r)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,pcols", [(1, 8192), (64, 256), (300, 1024)])
def test_reduce_scatter_flat_and_last_axis(emu, dt, rows, pcols):
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    nr, heap, comms = emu
    es = 4 if dt == torch.float32 else 2
    src = heap.alloc(rows * pcols * nr * es)
    d0, d1 = heap.alloc(rows * pcols * es), heap.alloc(rows * pcols * es)
    g = torch.Generator(device="cuda").manual_seed(1)
    for r in range(nr):
        heap.view(src, (rows, pcols * nr), dt, rank=r).copy_(torch.randn(rows, pcols * nr, device="cuda", generator=g).to(dt))
    cc = CO.cc_dtype(dt)
    for op in (N.SUM, N.MIN):
        CO.reduce_scatter(comms, cc, op, src, d0, rows, pcols, pcols * nr * es, pcols * es, use_mc=False, tight=False)
        CO.reduce_scatter(comms, cc, op, src, d1, rows, pcols, pcols * nr * es, pcols * es, use_mc=False, tight=True)
        torch.cuda.synchronize()
        total = None
        for r in range(nr):
            assert torch.equal(heap.view(d0, (rows, pcols), dt, rank=r), heap.view(d1, (rows, pcols), dt, rank=r)), (op, r)
        if op == N.SUM and dt == torch.float32:
            total = sum(heap.view(src, (rows, pcols * nr), dt, rank=r) for r in range(nr))
            for r in range(nr):
                torch.testing.assert_close(heap.view(d1, (rows, pcols), dt, rank=r), total[:, r * pcols:(r + 1) * pcols], rtol=1e-5, atol=1e-5)
