"""2-D (data x tensor-model parallel) process grid and dataset sharding.

Behavioural contract (reference model/func_impl.py:5-74 and
data/data_parallel_preprocess.py:3-59):

* model-parallel-major rank map: ``mp_idx = rank % mp_size``, ``dp_idx = rank // mp_size``
  (figs/group.png);
* ``mp_comm``: the ranks of one DP replica (world ranks ``dp_idx*mp_size + 0..mp_size-1``),
  ``dp_comm``: the ranks holding the same weight shard (stride ``mp_size``);
* column-parallel layers (``fc_q/fc_k/fc_v``, and ``fc1`` as an extension) split ``out_dim``;
  row-parallel layers (``fc_o``, and ``fc2``) split ``in_dim``;
* ``split_data`` hands every MP rank of a replica the same contiguous 1/dp slice, unshuffled.

On the device plane a sub-communicator is a subset of the world's peer-pointer table with its
own signal pad and (when NVLS is present) its own multicast object; nothing is copied when a
communicator is split, and splitting the same way twice returns the same communicator.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Tuple

COLUMN_PARALLEL = ("fc_q", "fc_k", "fc_v", "fc1")
ROW_PARALLEL = ("fc_o", "fc2")


@dataclass(frozen=True)
class GridPos:
    rank: int
    mp_size: int
    dp_size: int

    @property
    def mp_idx(self) -> int:
        return self.rank % self.mp_size

    @property
    def dp_idx(self) -> int:
        return self.rank // self.mp_size


def partition_dims(fc_layer: str, in_dim: int, out_dim: int, mp_size: int) -> Tuple[int, int]:
    """``(part_in_dim, part_out_dim)`` of one rank's weight shard."""
    if fc_layer in COLUMN_PARALLEL:
        return in_dim, out_dim // mp_size
    if fc_layer in ROW_PARALLEL:
        return in_dim // mp_size, out_dim
    raise ValueError(f"Invalid fc_layer: {fc_layer}.")


def grid_comms(comm: Any, rank: int, mp_size: int):
    """Split ``comm`` into this rank's model-parallel and data-parallel communicators.
    ``comm`` may be an ``MPI.Comm`` or a ``Communicator``: both accept ``Split`` by keyword."""
    pos = GridPos(rank, mp_size, 0)
    mp_comm = comm.Split(color=pos.dp_idx, key=pos.mp_idx)
    dp_comm = comm.Split(color=pos.mp_idx, key=pos.dp_idx)
    return pos.mp_idx, pos.dp_idx, mp_comm, dp_comm


def get_info(comm, rank: int, mp_size: int, dp_size: int, fc_layer: str, in_dim: int, out_dim: int):
    """Everything a rank needs to know about its place in the grid for one FC layer.

    Returns ``(mp_idx, dp_idx, mp_comm, dp_comm, part_in_dim, part_out_dim)``. The layer name
    is validated before any communication so that a bad name cannot leave peers inside a
    collective ``Split``."""
    part_in, part_out = partition_dims(fc_layer, in_dim, out_dim, mp_size)
    mp_idx, dp_idx, mp_comm, dp_comm = grid_comms(comm, rank, mp_size)
    return mp_idx, dp_idx, mp_comm, dp_comm, part_in, part_out


def split_data(x_train, y_train, mp_size: int, dp_size: int, rank: int):
    """This rank's share of the dataset: slice ``rank // mp_size`` of ``dp_size`` equal,
    contiguous, unshuffled leading-axis slices. Views, not copies -- works for NumPy arrays
    and torch tensors of any trailing shape."""
    per_replica = x_train.shape[0] // dp_size
    first = (rank // mp_size) * per_replica
    window = slice(first, first + per_replica)
    return x_train[window], y_train[window]
