"""A master / worker task farm on the ``mpi4py`` front end: rank 0 hands out work items, workers return results in
whatever order they finish (``Recv`` from ``MPI.ANY_SOURCE`` with a ``Status``), rank 0 answers the sender with the next
item or a stop tag. Nothing here is specific to this framework -- it is the program one would write against
mpi4py + ``mpirun``; it runs unchanged on the in-repo front end:

    python -m collective_communication_mpi_b200.launch -n 5 python examples/master_worker.py --items 40
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from mpi4py import MPI  # noqa: E402

WORK, STOP = 1, 2


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--items", type=int, default=40)
    args = ap.parse_args()
    comm = MPI.COMM_WORLD
    rank, n = comm.Get_rank(), comm.Get_size()
    if n < 2:
        raise SystemExit("needs at least one worker (run with -n 2 or more)")
    if rank == 0:
        todo = list(range(args.items))
        results, per_worker = {}, [0] * n
        active = 0
        for w in range(1, n):                                   # prime every worker (or stop it if there is no work)
            if todo:
                comm.send(todo.pop(), dest=w, tag=WORK)
                active += 1
            else:
                comm.send(None, dest=w, tag=STOP)
        st = MPI.Status()
        buf = np.empty(2, dtype=np.float64)
        while active:
            comm.Recv(buf, source=MPI.ANY_SOURCE, tag=MPI.ANY_TAG, status=st)
            w = st.Get_source()
            results[int(buf[0])] = buf[1]
            per_worker[w] += 1
            if todo:
                comm.send(todo.pop(), dest=w, tag=WORK)
            else:
                comm.send(None, dest=w, tag=STOP)
                active -= 1
        assert sorted(results) == list(range(args.items))
        assert all(abs(results[i] - sum(k * k for k in range(i + 1))) < 1e-9 for i in results)
        print(f"{args.items} items done by {n - 1} workers; items per worker: {per_worker[1:]}")
        print("FARM OK")
    else:
        st = MPI.Status()
        while True:
            item = comm.recv(source=0, tag=MPI.ANY_TAG, status=st)
            if st.Get_tag() == STOP:
                break
            comm.Send(np.array([item, sum(k * k for k in range(item + 1))], dtype=np.float64), dest=0, tag=item % 7)
    comm.Barrier()


if __name__ == "__main__":
    main()
