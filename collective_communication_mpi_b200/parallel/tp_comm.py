"""The four communication steps of the "naive" tensor-parallel FC layer.

Reference dataflow (model/func_impl.py:76-187; the GEMMs between the steps are not in the
reference at all)::

    fwd:  x_local (b,s,in/mp)  --gather last axis-->  x_full (b,s,in)          :76-91
          y_local = x_full @ W_local
          y_full  = gather last axis (y_local)                                   :94-109
    bwd:  dy_local = dy_full[..., mp_idx-th slice]                               :111-147
          dx_part  = dy_local @ W_local^T
          dx_local = reduce-scatter last axis (dx_part)                          :150-187

Each step here is ONE strided collective (kernels in csrc/collectives.cu, or the shared-memory
transport without GPUs) instead of pickle + ``np.concatenate`` / ``np.split`` + ``np.sum``.
Inputs may be NumPy arrays or torch tensors (CPU or CUDA); the result has the input's type and
dtype. A foreign communicator (e.g. a genuine mpi4py ``Comm``) gets the reference's own
object-collective recipe, so these functions stay drop-in.

The fused versions -- gather folded into the consuming GEMM, reduce-scatter folded into the
producing GEMM's epilogue -- live in ``ops/fused.py``; these standalone steps are the API the
reference exposes and the oracle the fused kernels are tested against.
"""
from __future__ import annotations

import numpy as np


def _native_comm(comm):
    """Our communicator (or the byte-counting wrapper around it) exposes the strided ops."""
    return hasattr(comm, "gather_last_axis") and hasattr(comm, "reduce_scatter_last_axis")


def _is_tensor(x) -> bool:
    return type(x).__module__.startswith("torch")


def _cat_last(parts):
    if _is_tensor(parts[0]):
        import torch
        return torch.cat(list(parts), dim=-1)
    return np.concatenate(parts, axis=-1)


def collect_forward_input(x, mp_comm, mp_size: int):
    """``(b, s, part_in)`` on every MP rank -> ``(b, s, part_in * mp)`` on all of them."""
    if _native_comm(mp_comm):
        return mp_comm.gather_last_axis(x)
    return _cat_last(mp_comm.allgather(x))


def collect_forward_output(out, mp_comm, mp_size: int):
    """``(b, s, part_out)`` -> ``(b, s, part_out * mp)``; same mechanism on the layer output."""
    if _native_comm(mp_comm):
        return mp_comm.gather_last_axis(out)
    return _cat_last(mp_comm.allgather(out))


def collect_backward_output(output_grad, mp_group_idx: int, mp_size: int):
    """This rank's slice of the full output gradient: no communication, a strided view."""
    width = output_grad.shape[-1] // mp_size
    return output_grad[..., mp_group_idx * width:(mp_group_idx + 1) * width]


def collect_backward_x(grad_x, mp_comm, mp_size: int):
    """Sum the ranks' partial ``grad_x`` and keep this rank's 1/mp of the last axis."""
    if _native_comm(mp_comm):
        return mp_comm.reduce_scatter_last_axis(grad_x)
    if _is_tensor(grad_x):
        import torch
        pieces = mp_comm.alltoall(list(torch.chunk(grad_x, mp_size, dim=-1)))
        return torch.stack(list(pieces)).sum(0)
    pieces = mp_comm.alltoall(np.split(grad_x, mp_size, axis=-1))
    return np.sum(pieces, axis=0)
