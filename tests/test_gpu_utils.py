"""Device-side utilities: input prefetcher / async loss reader, launch counter, symmetric heap."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def test_prefetcher_delivers_batches_in_order_and_reader_returns_every_value():
    from collective_communication_mpi_b200.utils.data import AsyncScalarReader, DevicePrefetcher
    host = [[torch.full((1 << 16,), float(i)).pin_memory(), torch.full((8,), float(-i)).pin_memory()] for i in range(6)]
    pf, rd = DevicePrefetcher(host[0]), AsyncScalarReader()
    got = []
    pf.submit(host[0])
    for i in range(6):
        a, b = pf.get()
        if i + 1 < 6:
            pf.submit(host[i + 1])
        rd.push(a.sum() / a.numel() + b.sum() / b.numel() * 0)   # "loss" of this step
        pf.release()
        if rd.pending > 2:
            got.append(rd.pop())
    while rd.pending:
        got.append(rd.pop())
    assert got == [float(i) for i in range(6)]
    assert pf.bytes_per_batch == (1 << 16) * 4 + 32


def test_launch_counter_counts_our_kernels():
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops.gemm import gemm
    a = torch.randn(128, 64, device="cuda").bfloat16()
    before = N.LAUNCHES[0]
    gemm(a, a)
    gemm(a, a)
    assert N.LAUNCHES[0] - before == 2


def test_symmetric_heap_allocator_reuses_and_merges_blocks():
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap
    heap = SymmetricHeap.create_emulated(1, 8 << 20)
    try:
        a, b, c = heap.alloc(1 << 20), heap.alloc(1 << 20), heap.alloc(1 << 20)
        heap.free(a); heap.free(b)
        d = heap.alloc(2 << 20)          # only fits if the two freed neighbours were merged
        assert d == a
        with pytest.raises(MemoryError):
            heap.alloc(64 << 20)
        t = heap.view(c, (4, 8), torch.float32)
        t.fill_(3.0)
        assert heap.offset_of(t) == c and float(heap.view(c, (32,), torch.float32).sum()) == 96.0
    finally:
        heap.close()
