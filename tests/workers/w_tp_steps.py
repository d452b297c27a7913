"""Rank-side body of the 2-D layout / tensor-parallel communication-step tests.

Scenarios follow the reference's MPI tests (tests/test_get_info.py: 8 ranks, mp=4 x dp=2, an
(8,10) int64 arange whose rows are all-reduced over mp_comm and dp_comm;
tests/test_transformer_forward.py: last-axis gather of (4,8,2) float64 shards of a (4,8,8)
arange; tests/test_transformer_backward.py: slice of (1,4,8), reduce-scatter of per-rank
(1,3,8) into (1,3,2)), with the expected values computed here in closed form.
``--cuda`` repeats them on CUDA tensors."""
import sys

import numpy as np

from mpi4py import MPI
from model.func_impl import (get_info, naive_collect_backward_output, naive_collect_backward_x,
                             naive_collect_forward_input, naive_collect_forward_output)
from mpi_wrapper import Communicator

CUDA = "--cuda" in sys.argv
comm = MPI.COMM_WORLD
rank, n = comm.Get_rank(), comm.Get_size()
if CUDA:
    import torch
    torch.cuda.set_device(rank)


def put(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda() if CUDA else a


def get(a):
    return a.cpu().numpy() if CUDA else a


def grid_case(mp_size, dp_size, layer, in_dim, out_dim, want_in, want_out, wrap):
    base = Communicator(comm) if wrap else comm
    mp_idx, dp_idx, mp_comm, dp_comm, pin, pout = get_info(
        comm=base, rank=rank, mp_size=mp_size, dp_size=dp_size, fc_layer=layer, in_dim=in_dim, out_dim=out_dim)
    assert (mp_idx, dp_idx) == (rank % mp_size, rank // mp_size)
    assert (pin, pout) == (want_in, want_out)
    assert mp_comm.Get_size() == mp_size and dp_comm.Get_size() == dp_size
    assert mp_comm.Get_rank() == mp_idx and dp_comm.Get_rank() == dp_idx
    table = np.arange(n * 10).reshape(n, 10)
    row = put(table[rank].copy())
    got_mp, got_dp = put(np.empty(10, dtype=table.dtype)), put(np.empty(10, dtype=table.dtype))
    mp_comm.Allreduce(row, got_mp, op=MPI.SUM)
    dp_comm.Allreduce(row, got_dp, op=MPI.SUM)
    replica = [r for r in range(n) if r // mp_size == dp_idx]
    shard_peers = [r for r in range(n) if r % mp_size == mp_idx]
    np.testing.assert_allclose(get(got_mp), table[replica].sum(0))
    np.testing.assert_allclose(get(got_dp), table[shard_peers].sum(0))


if n == 8:
    for wrap in (False, True):
        grid_case(4, 2, "fc_q", 768, 256, 768, 64, wrap)
        grid_case(4, 2, "fc_o", 256, 10, 64, 10, wrap)
        grid_case(2, 4, "fc1", 64, 256, 64, 128, wrap)
        grid_case(8, 1, "fc2", 256, 64, 32, 64, wrap)
        grid_case(1, 8, "fc_k", 16, 16, 16, 16, wrap)
    try:
        get_info(comm=comm, rank=rank, mp_size=4, dp_size=2, fc_layer="fc_bad", in_dim=8, out_dim=8)
        raise AssertionError("bad layer name must raise")
    except ValueError as e:
        assert "Invalid fc_layer: fc_bad." in str(e)

# ---- forward gathers on the last axis (mp group = first min(n,4) ranks' worth of columns) ------
mp = 4
full = np.arange(4 * 8 * 8, dtype=np.float64).reshape(4, 8, 8)
shard = full[:, :, rank * 2:(rank + 1) * 2]            # non-contiguous view; zero-width for rank >= 4
for collect in (naive_collect_forward_input, naive_collect_forward_output):
    kw = {"x": put(shard)} if collect is naive_collect_forward_input else {"out": put(shard)}
    got = collect(mp_comm=comm, mp_size=mp, **kw)
    assert get(got).dtype == full.dtype
    np.testing.assert_allclose(get(got), full[:, :, :2 * n] if n < 4 else full)

if n == 4:
    for dtype in (np.float32, np.int64):
        wide = (np.arange(2 * 5 * 12).reshape(2, 5, 12) % 17).astype(dtype)
        got = naive_collect_forward_input(x=put(wide[..., rank * 3:(rank + 1) * 3]), mp_comm=Communicator(comm), mp_size=4)
        np.testing.assert_array_equal(get(got), wide)
        assert get(got).dtype == dtype

    # ---- backward: slice, then reduce-scatter on the last axis -------------------------------
    dy = np.arange(1 * 4 * 8, dtype=np.float64).reshape(1, 4, 8)
    piece = naive_collect_backward_output(output_grad=put(dy), mp_group_idx=rank, mp_size=4)
    assert get(piece).dtype == dy.dtype
    np.testing.assert_allclose(get(piece), dy[:, :, rank * 2:(rank + 1) * 2])

    stack = np.arange(4 * 3 * 8, dtype=np.float64).reshape(4, 3, 8)
    got = naive_collect_backward_x(grad_x=put(stack[rank:rank + 1]), mp_comm=comm, mp_size=4)
    assert get(got).dtype == stack.dtype and get(got).shape == (1, 3, 2)
    np.testing.assert_allclose(get(got), stack.sum(0, keepdims=True)[:, :, rank * 2:(rank + 1) * 2])

    rng = np.random.RandomState(0)
    parts = rng.randn(4, 2, 16, 64).astype(np.float32)
    got = naive_collect_backward_x(grad_x=put(parts[rank]), mp_comm=Communicator(comm), mp_size=4)
    np.testing.assert_allclose(get(got), parts.sum(0)[..., rank * 16:(rank + 1) * 16], rtol=1e-5, atol=1e-5)

comm.Barrier()
if rank == 0:
    print(f"TP STEPS OK ranks={n} cuda={CUDA}")
