"""CUDA-graph replay of stand-alone collectives (device-side epochs) with emulated ranks on one GPU."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def test_a_captured_chain_of_small_allreduces_replays_correctly():
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    from collective_communication_mpi_b200.utils.graphs import capture
    nr, cnt = 2, 1024
    heap = SymmetricHeap.create_emulated(nr, 16 << 20)
    try:
        comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
        a, b, t = heap.alloc(cnt * 4), heap.alloc(cnt * 4), heap.alloc(cnt * 4)
        views = [heap.view(a, (cnt,), torch.float32, rank=r) for r in range(nr)]

        def chain():                      # a -> b -> a -> ... : 10 dependent all-reduces, each doubles the sum
            for i in range(10):
                src, dst = (a, b) if i % 2 == 0 else (b, a)
                CO.allreduce(comms, N.F32, N.SUM, src, dst, cnt, tmp_off=t, algo=N.AR_ONESHOT)

        for r in range(nr):
            views[r].fill_(1.0)
        step = capture(chain)             # warm-up calls already multiplied the values; reset before replays
        for rep in range(3):
            for r in range(nr):
                views[r].fill_(float(r + 1))
            torch.cuda.synchronize()
            step()
            torch.cuda.synchronize()
            want = float(sum(range(1, nr + 1))) * (nr ** 9)      # first reduce sums 1..n, the other 9 multiply by n
            for r in range(nr):
                assert torch.all(views[r] == want), (rep, r, views[r][:4].tolist(), want)
    finally:
        heap.close()
