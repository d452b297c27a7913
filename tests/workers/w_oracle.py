"""Rank-side: random-data collectives and a full tensor-parallel FC layer (NumPy GEMMs around the
func_impl communication steps) checked against the all-ranks-at-once oracles of utils/oracle.py,
plus per-op statistics of the Communicator."""
import numpy as np

from mpi4py import MPI
from model.func_impl import (get_info, naive_collect_backward_output, naive_collect_backward_x,
                             naive_collect_forward_input, naive_collect_forward_output)
from mpi_wrapper import Communicator
from collective_communication_mpi_b200.utils import oracle as O

world = MPI.COMM_WORLD
rank, n = world.Get_rank(), world.Get_size()
rng = np.random.default_rng(7)                       # same stream on every rank: everyone can build all inputs

c = Communicator(world)
for dtype in (np.float64, np.float32, np.int64):
    for count in (n, 16 * n, 1000 * n):
        parts = [(rng.standard_normal(count) * 10).astype(dtype) for _ in range(n)]
        for op, name in ((MPI.SUM, "sum"), (MPI.MIN, "min"), (MPI.MAX, "max")):
            out = np.empty(count, dtype=dtype)
            c.Allreduce(parts[rank], out, op=op)
            np.testing.assert_allclose(out, O.allreduce(parts, name), rtol=1e-5)
            mine = np.empty(count, dtype=dtype)
            c.myAllreduce(parts[rank], mine, op=op)
            np.testing.assert_array_equal(mine, out)
        out = np.empty(count * n, dtype=dtype)
        c.Allgather(parts[rank], out)
        np.testing.assert_array_equal(out, O.allgather(parts))
        out = np.empty(count // n, dtype=dtype)
        c.Reduce_scatter(parts[rank], out)
        np.testing.assert_allclose(out, O.reduce_scatter(parts)[rank], rtol=1e-5)
        out = np.empty(count, dtype=dtype)
        c.Alltoall(parts[rank], out)
        np.testing.assert_array_equal(out, O.alltoall(parts)[rank])
        out2 = np.empty(count, dtype=dtype)
        c.myAlltoall(parts[rank], out2)
        np.testing.assert_array_equal(out2, out)
rep = c.stats.report()
assert rep["Allreduce"]["calls"] == 27 and rep["Allgather"]["calls"] == 9 and rep["Alltoall"]["bytes"] > 0
assert all(v["seconds"] > 0 for v in rep.values())

# ---- one column-parallel and one row-parallel FC layer, forward and backward -----------------
mp = n
mp_idx, dp_idx, mp_comm, dp_comm, pin, pout = get_info(comm=world, rank=rank, mp_size=mp, dp_size=1, fc_layer="fc_q",
                                                       in_dim=8 * mp, out_dim=12 * mp)
b, s, din, dout = 2, 5, 8 * mp, 12 * mp
xs = [rng.standard_normal((b, s, din // mp)) for _ in range(mp)]
ws = [rng.standard_normal((din, dout // mp)) for _ in range(mp)]
dys = [rng.standard_normal((b, s, dout // mp)) for _ in range(mp)]
want = O.column_parallel_fc([x.reshape(-1, din // mp) for x in xs], ws, [d.reshape(-1, dout // mp) for d in dys])[rank]
x_full = naive_collect_forward_input(x=xs[rank], mp_comm=mp_comm, mp_size=mp)
y_loc = x_full @ ws[rank]
y_full = naive_collect_forward_output(out=y_loc, mp_comm=mp_comm, mp_size=mp)
np.testing.assert_allclose(y_loc.reshape(-1, dout // mp), want[0])
np.testing.assert_allclose(y_full, np.concatenate([np.concatenate(xs, -1) @ w for w in ws], -1))
dy_full = np.concatenate(dys, -1)
dy_loc = naive_collect_backward_output(output_grad=dy_full, mp_group_idx=mp_idx, mp_size=mp)
dw = np.tensordot(x_full, dy_loc, axes=([0, 1], [0, 1]))
dx_loc = naive_collect_backward_x(grad_x=dy_loc @ ws[rank].T, mp_comm=mp_comm, mp_size=mp)
np.testing.assert_allclose(dw, want[2])
np.testing.assert_allclose(dx_loc.reshape(-1, din // mp), want[1])

wr = [rng.standard_normal((din // mp, dout)) for _ in range(mp)]
want = O.row_parallel_fc([x.reshape(-1, din // mp) for x in xs], wr, [d.reshape(-1, dout // mp) for d in dys])[rank]
y_loc = naive_collect_backward_x(grad_x=xs[rank] @ wr[rank], mp_comm=mp_comm, mp_size=mp)       # reduce-scatter of the partial output
np.testing.assert_allclose(y_loc.reshape(-1, dout // mp), want[0])
dy_full = naive_collect_forward_input(x=dys[rank], mp_comm=mp_comm, mp_size=mp)
np.testing.assert_allclose((dy_full @ wr[rank].T).reshape(-1, din // mp), want[1])
np.testing.assert_allclose(np.tensordot(xs[rank], dy_full, axes=([0, 1], [0, 1])), want[2])
world.Barrier()
if rank == 0:
    print(f"ORACLE OK ranks={n}")
