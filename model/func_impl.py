"""Reference-compatible entry points (same names, same keyword signatures as the reference's
model/func_impl.py:5,76,94,111,150); the implementations live in
``collective_communication_mpi_b200.parallel``.

``fc1`` / ``fc2`` are accepted in addition to ``fc_q/fc_k/fc_v/fc_o`` (column- and row-parallel
respectively); any other layer name raises ``ValueError`` like the reference (:71-72).
"""
import numpy as np  # noqa: F401  (kept importable for users that reach for func_impl.np)

from collective_communication_mpi_b200.parallel import layout as _layout
from collective_communication_mpi_b200.parallel import tp_comm as _tp


def get_info(comm, rank: int, mp_size: int, dp_size: int, fc_layer: str, in_dim: int, out_dim: int):
    return _layout.get_info(comm=comm, rank=rank, mp_size=mp_size, dp_size=dp_size, fc_layer=fc_layer,
                            in_dim=in_dim, out_dim=out_dim)


def naive_collect_forward_input(x, mp_comm, mp_size: int):
    return _tp.collect_forward_input(x, mp_comm, mp_size)


def naive_collect_forward_output(out, mp_comm, mp_size: int):
    return _tp.collect_forward_output(out, mp_comm, mp_size)


def naive_collect_backward_output(output_grad, mp_group_idx: int, mp_size: int):
    return _tp.collect_backward_output(output_grad, mp_group_idx, mp_size)


def naive_collect_backward_x(grad_x, mp_comm, mp_size: int):
    return _tp.collect_backward_x(grad_x, mp_comm, mp_size)


get_info.__doc__ = _layout.get_info.__doc__
naive_collect_forward_input.__doc__ = _tp.collect_forward_input.__doc__
naive_collect_forward_output.__doc__ = _tp.collect_forward_output.__doc__
naive_collect_backward_output.__doc__ = _tp.collect_backward_output.__doc__
naive_collect_backward_x.__doc__ = _tp.collect_backward_x.__doc__
