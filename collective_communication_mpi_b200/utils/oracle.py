"""NumPy oracles: closed-form restatements of every collective and of the tensor-parallel FC
layer's forward / backward, computed from ALL ranks' inputs at once.

They restate what MPI defines (and what the reference gets from it: comm.py:16-61) and the
dataflow of func_impl.py:76-187; tests compare both transports and the fused kernels against
them. ``parts`` is always the list of per-rank arrays in rank order.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

_FOLD = {"sum": np.add, "min": np.minimum, "max": np.maximum}


def allreduce(parts: Sequence[np.ndarray], op: str = "sum") -> np.ndarray:
    """Rank-ordered fold (the order of the reference's root loop, comm.py:85-93)."""
    acc = np.array(parts[0], copy=True)
    for p in parts[1:]:
        _FOLD[op](acc, p, out=acc)
    return acc


def allgather(parts: Sequence[np.ndarray]) -> np.ndarray:
    return np.concatenate([np.ravel(p) for p in parts])


def reduce_scatter(parts: Sequence[np.ndarray], op: str = "sum") -> List[np.ndarray]:
    n = len(parts)
    full = allreduce([np.ravel(p) for p in parts], op)
    return list(np.split(full, n))


def alltoall(parts: Sequence[np.ndarray]) -> List[np.ndarray]:
    n = len(parts)
    blocks = [np.split(np.ravel(p), n) for p in parts]
    return [np.concatenate([blocks[src][dst] for src in range(n)]) for dst in range(n)]


def gather_last_axis(parts: Sequence[np.ndarray]) -> np.ndarray:
    """func_impl.py:89-90 / :107-108."""
    return np.concatenate(list(parts), axis=-1)


def reduce_scatter_last_axis(parts: Sequence[np.ndarray]) -> List[np.ndarray]:
    """func_impl.py:182-186: every rank's slice of the rank-ordered sum."""
    total = allreduce(list(parts), "sum")
    return list(np.split(total, len(parts), axis=-1))


def bytes_accounted(op: str, src_bytes: int, dst_bytes: int, n: int, rank: int = 0) -> int:
    """``Communicator.total_bytes_transferred`` increment of one call."""
    if op == "Allreduce":
        return src_bytes * 2 * (n - 1)
    if op in ("Allgather", "Reduce_scatter"):
        return (src_bytes + dst_bytes) * (n - 1)
    if op == "Alltoall":
        return (src_bytes // n + dst_bytes // n) * (n - 1)
    if op == "myAllreduce":
        return 2 * src_bytes * (n - 1) if rank == 0 else 2 * src_bytes
    if op in ("myAlltoall", "myAlltoall2"):
        return 2 * (src_bytes // n) * (n - 1)
    raise ValueError(op)


def column_parallel_fc(x_shards, w_shards, dy_shards):
    """Column-parallel FC layer over ``mp`` ranks. ``x_shards[r]``: (M, in/mp), ``w_shards[r]``:
    (in, out/mp), ``dy_shards[r]``: (M, out/mp). Returns per-rank ``(y, dx, dw)``."""
    x_full = gather_last_axis(x_shards)
    ys = [x_full @ w for w in w_shards]
    dws = [x_full.T @ dy for dy in dy_shards]
    dxs = reduce_scatter_last_axis([dy @ w.T for dy, w in zip(dy_shards, w_shards)])
    return list(zip(ys, dxs, dws))


def row_parallel_fc(x_shards, w_shards, dy_shards):
    """Row-parallel FC layer. ``x_shards[r]``: (M, in/mp), ``w_shards[r]``: (in/mp, out),
    ``dy_shards[r]``: (M, out/mp) (the output is reduce-scattered on its last axis)."""
    ys = reduce_scatter_last_axis([x @ w for x, w in zip(x_shards, w_shards)])
    dy_full = gather_last_axis(dy_shards)
    dxs = [dy_full @ w.T for w in w_shards]
    dws = [x.T @ dy_full for x in x_shards]
    return list(zip(ys, dxs, dws))
