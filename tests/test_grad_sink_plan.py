"""Bucket plan of the data-parallel gradient sink (parallel/dp.py), on CPU tensors: what goes where in the
flat buffer, in which order buckets fill up, when a bucket is handed to the reduction."""
import torch

from collective_communication_mpi_b200.parallel.dp import GradSink


class _Comm:
    """Stand-in for a data-parallel communicator of ``size`` replicas that all hold this rank's data."""

    def __init__(self, size):
        self._size, self.calls = size, []

    def Get_size(self):
        return self._size

    def Allreduce(self, src, dst):
        self.calls.append(src.numel())
        dst.copy_(src.float().mul(self._size).to(dst.dtype))       # every replica contributes the same values


class _Layer:
    def __init__(self, rows, cols):
        self.weight = torch.nn.Parameter(torch.zeros(rows, cols, dtype=torch.bfloat16))
        self.mp = 1


def test_buckets_follow_backward_order_and_views_alias_the_flat_buffer():
    layers = [_Layer(64, 100), _Layer(100, 33), _Layer(7, 9), _Layer(512, 1024)]
    sink = GradSink(_Comm(1), layers, bucket_mb=1)
    pad = lambda n: (n + 127) // 128 * 128                          # noqa: E731
    assert sink.flat.numel() == sum(pad(l.weight.numel()) for l in layers)
    # the last layer's gradient is produced first in backward, so it sits first in the buffer
    off = 0
    for l in reversed(layers):
        v = sink.view_for(l.weight)
        assert v.shape == l.weight.shape and v.data_ptr() == sink.flat.data_ptr() + 2 * off
        assert (2 * off) % 256 == 0, "every gradient view is 256-byte aligned (vector loads of the reduction)"
        off += pad(l.weight.numel())
        assert l.grad_sink is sink
    # 1 MB buckets: the 512 x 1024 bf16 gradient (1 MB) fills the first bucket alone, the three small ones share the next
    assert [(b["lo"], b["hi"], b["want"]) for b in sink.buckets] == [(0, 512 * 1024, 1), (512 * 1024, sink.flat.numel(), 3)]
    v = sink.view_for(layers[2].weight)
    v.fill_(3.0)
    lo = pad(512 * 1024) + 0
    assert float(sink.flat[lo:lo + 63].float().sum()) == 3.0 * 63   # writes through the view land in the flat buffer


def test_a_bucket_is_reduced_once_when_its_last_gradient_arrives_and_is_averaged():
    layers = [_Layer(64, 64), _Layer(64, 64), _Layer(1024, 1024)]
    comm = _Comm(4)
    sink = GradSink(comm, layers, bucket_mb=1, overlap=True)
    assert not sink.overlap and not sink.fused_update and not sink.partitioned      # CPU: no streams, no NVLS
    assert len(sink.buckets) == 2
    for l in layers:
        sink.view_for(l.weight).fill_(2.0)
    sink.ready(layers[2].weight)                                     # backward order: last layer first
    assert comm.calls == [1024 * 1024] and sink.launches == 1
    sink.ready(layers[1].weight)
    assert comm.calls == [1024 * 1024], "bucket of two small layers waits for both"
    sink.ready(layers[0].weight)
    assert comm.calls == [1024 * 1024, 2 * 64 * 64] and sink.launches == 2
    for l in layers:                                                 # sum over 4 replicas x 1/4 = the replica's own value
        assert torch.all(sink.view_for(l.weight) == 2.0)
    # next step: counters start over
    sink.ready(layers[2].weight)
    assert sink.launches == 3


def test_single_replica_reduces_nothing():
    layers = [_Layer(8, 8)]
    comm = _Comm(1)
    sink = GradSink(comm, layers)
    sink.ready(layers[0].weight)
    assert comm.calls == [] and sink.launches == 0
    assert sink.grads()[0].data_ptr() == sink.flat.data_ptr()
