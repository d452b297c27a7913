"""Differential test against the UNMODIFIED reference: its ``Communicator`` (mpi_wrapper/comm.py), its
``model/func_impl.py`` and its ``split_data`` are loaded from ``baseline/_ref`` under private module names and
run side by side with this repository's, on the same communicators, for seeded random shapes, dtypes and
operators. Results must be equal bit for bit and ``total_bytes_transferred`` must advance identically."""
import importlib.util
import os
import sys

import numpy as np

from mpi4py import MPI

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "baseline", "_ref")


def load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


ref_comm = load("_ref_comm", "mpi_wrapper/comm.py")
ref_func = load("_ref_func", "model/func_impl.py")
ref_data = load("_ref_data", "data/data_parallel_preprocess.py")
from mpi_wrapper import Communicator                                   # noqa: E402  (ours)
from model import func_impl as our_func                                # noqa: E402
from data.data_parallel_preprocess import split_data as our_split      # noqa: E402

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
ours, theirs = Communicator(world), ref_comm.Communicator(world)
rng = np.random.default_rng(2024)                                       # same stream on every rank: same shapes
mine = np.random.default_rng(1000 + rank)                               # per-rank payloads
cases = 0


def same_bytes(what):
    assert ours.total_bytes_transferred == theirs.total_bytes_transferred, \
        f"{what}: accounting differs: ours {ours.total_bytes_transferred}, reference {theirs.total_bytes_transferred}"


for trial in range(40):
    dt = [np.int64, np.int32, np.float64, np.float32][trial % 4]
    per = int(rng.integers(1, 9))
    count = per * n
    payload = (mine.integers(-50, 50, size=count)).astype(dt)
    op = [MPI.SUM, MPI.MIN, MPI.MAX][trial % 3]
    # library collectives through the byte-counting wrapper
    a, b = np.empty_like(payload), np.empty_like(payload)
    ours.Allreduce(payload, a, op)
    theirs.Allreduce(payload, b, op)
    assert np.array_equal(a, b)
    same_bytes("Allreduce")
    a, b = np.empty(count * n, dtype=dt), np.empty(count * n, dtype=dt)
    ours.Allgather(payload, a)
    theirs.Allgather(payload, b)
    assert np.array_equal(a, b)
    same_bytes("Allgather")
    a, b = np.empty(per, dtype=dt), np.empty(per, dtype=dt)
    ours.Reduce_scatter(payload, a, op)
    theirs.Reduce_scatter(payload, b, op)
    assert np.array_equal(a, b)
    same_bytes("Reduce_scatter")
    a, b = np.empty_like(payload), np.empty_like(payload)
    ours.Alltoall(payload, a)
    theirs.Alltoall(payload, b)
    assert np.array_equal(a, b)
    same_bytes("Alltoall")
    # the hand-rolled collectives (comm.py:63-199)
    a, b = np.empty_like(payload), np.empty_like(payload)
    ours.myAllreduce(payload, a, op)
    theirs.myAllreduce(payload, b, op)
    assert np.array_equal(a, b), ("myAllreduce", dt, op)
    same_bytes("myAllreduce")
    a, b = np.empty_like(payload), np.empty_like(payload)
    ours.myAlltoall(payload, a)
    theirs.myAlltoall(payload, b)
    assert np.array_equal(a, b)
    same_bytes("myAlltoall")
    cases += 6

# get_info on every factorisation of the world, every layer name the reference knows (on the raw communicator, as
# the reference's tests call it: its byte-counting wrapper has no lowercase allgather / alltoall)
for mp in [m for m in range(1, n + 1) if n % m == 0]:
    dp = n // mp
    for layer in ("fc_q", "fc_k", "fc_v", "fc_o"):
        in_dim, out_dim = 16 * mp * int(rng.integers(1, 4)), 8 * mp * int(rng.integers(1, 4))
        o = our_func.get_info(comm=world, rank=rank, mp_size=mp, dp_size=dp, fc_layer=layer, in_dim=in_dim, out_dim=out_dim)
        t = ref_func.get_info(comm=world, rank=rank, mp_size=mp, dp_size=dp, fc_layer=layer, in_dim=in_dim, out_dim=out_dim)
        assert (o[0], o[1], o[4], o[5]) == (t[0], t[1], t[4], t[5]), (layer, mp, dp, o, t)
        assert (o[2].Get_size(), o[2].Get_rank(), o[3].Get_size(), o[3].Get_rank()) == \
               (t[2].Get_size(), t[2].Get_rank(), t[3].Get_size(), t[3].Get_rank())
        cases += 1
        if layer != "fc_q":
            continue
        # the four communication steps of the layer on this grid's model-parallel communicator
        mp_o, mp_t = o[2], t[2]
        bsz, seq, part = int(rng.integers(1, 4)), int(rng.integers(1, 6)), int(rng.integers(1, 5))
        x = mine.standard_normal((bsz, seq, part))
        assert np.array_equal(our_func.naive_collect_forward_input(x, mp_o, mp), ref_func.naive_collect_forward_input(x, mp_t, mp))
        assert np.array_equal(our_func.naive_collect_forward_output(x, mp_o, mp), ref_func.naive_collect_forward_output(x, mp_t, mp))
        gfull = mine.standard_normal((bsz, seq, part * mp))
        assert np.array_equal(our_func.naive_collect_backward_output(gfull, o[0], mp), ref_func.naive_collect_backward_output(gfull, t[0], mp))
        dx_o = our_func.naive_collect_backward_x(gfull, mp_o, mp)
        dx_t = ref_func.naive_collect_backward_x(gfull, mp_t, mp)
        assert dx_o.shape == dx_t.shape and dx_o.dtype == dx_t.dtype
        assert np.array_equal(dx_o, dx_t), "reduce-scatter: rank-ordered sum must reproduce np.sum(axis=0) bit for bit"
        cases += 4
    # split_data is rank arithmetic only (no communication): every rank checks every rank's shard
    N_ = mp * dp * int(rng.integers(1, 4))
    xs, ys = rng.standard_normal((N_, 3, 2)), rng.standard_normal((N_, 2))
    for r in range(n):
        ox, oy = our_split(xs, ys, mp_size=mp, dp_size=dp, rank=r)
        tx, ty = ref_data.split_data(xs, ys, mp_size=mp, dp_size=dp, rank=r)
        assert np.array_equal(ox, tx) and np.array_equal(oy, ty)
    cases += 1

world.Barrier()
if rank == 0:
    print(f"DIFFERENTIAL OK ranks={n} cases={cases}")
