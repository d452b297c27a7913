"""CUDA-graph capture of launch-bound sequences.

The stand-alone collectives keep their barrier epochs on the device (csrc/collectives.cu: every launch reads
and advances the epoch words of its communicator's signal pad), so a captured launch can be replayed any
number of times -- unlike the fused GEMM+collective kernels, whose epoch and buffer parity are host
arguments. Small-message collectives are launch-bound (11-13 us per call at 8 GPUs),
which is where replaying a graph of many calls pays: the reference's own benchmark is a loop of 100 tiny
all-reduces (mpi-test.py:40-98).

    step = capture(lambda: [CO.allreduce(comms, ...) for _ in range(100)])
    step()          # one graph launch instead of 100 kernel launches

Every rank must capture and replay the same sequence. Kernels go to ``torch.cuda.current_stream()``, which
is the capture stream inside ``torch.cuda.graph``; nothing in the wrappers allocates or synchronises.
Run on a B200 in round 2 (tests/test_gpu_graphs.py).
"""
from __future__ import annotations

from typing import Callable

import torch


def capture(fn: Callable[[], object], warmup: int = 2) -> Callable[[], object]:
    """Run ``fn`` ``warmup`` times on a side stream (first-use set-up such as function attributes must not
    happen during capture), capture one more call into a CUDA graph and return a callable that replays it
    and hands back what ``fn`` returned during capture (tensors are the graph's static outputs)."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(max(1, warmup)):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        result = fn()

    def replay():
        graph.replay()
        return result

    replay.graph = graph
    return replay
