"""The ``mpi-test.py`` demos and micro-benchmarks (reference mpi-test.py:6-241), table-driven.

    python -m collective_communication_mpi_b200.launch -n 8 python mpi-test.py --test_case myallreduce

Same seven ``--test_case`` choices, message shapes and printed lines as the reference. Additions
that do not change the defaults: ``--runs`` (timed repetitions, reference: 100), ``--count``
(elements per message, reference: 100 for the all-reduce cases) and ``--device`` (put the arrays
on this rank's GPU as torch tensors, so the timed region is the device kernel alone instead of
host staging + kernel).
"""
from __future__ import annotations

import argparse

import numpy as np

CASES = ["allreduce", "allgather", "reduce_scatter", "split", "alltoall", "myallreduce", "myalltoall"]


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser()
    p.add_argument("--test_case", type=str, default="", choices=CASES + [""],
                   help="MPI names for different toy examples")
    p.add_argument("--runs", type=int, default=100, help="timed repetitions of the my* comparisons")
    p.add_argument("--count", type=int, default=100, help="elements per rank in the all-reduce cases")
    p.add_argument("--device", action="store_true", help="keep the arrays on the GPU (torch tensors)")
    return p


class _Arrays:
    """Array factory: NumPy on the host (reference behaviour) or torch on this rank's GPU."""

    def __init__(self, device: bool, rank: int):
        self.device = device
        if device:
            import torch
            self.torch = torch
            self.dev = torch.device("cuda", torch.cuda.current_device())

    def randint(self, hi, n):
        a = np.random.randint(0, hi, n)
        return self.torch.from_numpy(a).to(self.dev) if self.device else a

    def empty(self, n):
        return self.torch.empty(n, dtype=self.torch.int64, device=self.dev) if self.device else np.empty(n, dtype=int)

    def array(self, values):
        a = np.array(values, dtype=int)
        return self.torch.from_numpy(a).to(self.dev) if self.device else a

    def show(self, a):
        return str(a.cpu().numpy() if self.device else a)

    def equal(self, a, b):
        return bool((a == b).all())

    def sync(self):
        if self.device:
            self.torch.cuda.synchronize()


def _timed(comm, MPI, fn, arrays):
    """Barrier-bracketed wall time of one call, as mpi-test.py:59-64 does it."""
    comm.Barrier()
    t0 = MPI.Wtime()
    fn()
    arrays.sync()
    comm.Barrier()
    return MPI.Wtime() - t0


def _compare(comm, MPI, args, arrays, make_inputs, run_lib, run_mine, lib_name, my_name):
    """The shared body of the ``myallreduce`` / ``myalltoall`` cases (mpi-test.py:40-98,178-239):
    time library vs hand-rolled, check equality every run, print the summary on rank 0."""
    rank = comm.Get_rank()
    t_lib, t_mine, all_ok = [], [], True
    for run in range(args.runs):
        src, out_lib, out_mine = make_inputs()
        t_lib.append(_timed(comm, MPI, lambda: run_lib(src, out_lib), arrays))
        t_mine.append(_timed(comm, MPI, lambda: run_mine(src, out_mine), arrays))
        if not arrays.equal(out_lib, out_mine):
            all_ok = False
            print("Rank {}: Run {}: ERROR: {} result does not match {}".format(rank, run, my_name, lib_name))
            if rank == 0:
                print("{} result:   ".format(lib_name), arrays.show(out_lib))
                print("{} result: ".format(my_name), arrays.show(out_mine))
        elif rank == 0:
            print("Run {}: Correct results.".format(run))
    if rank == 0:
        print("\nSummary over {} runs:".format(args.runs))
        print("All runs produced correct results." if all_ok else "Some runs produced incorrect results!")
        w = len("Average  time:") + max(len(lib_name), len(my_name)) + 1
        print(("Average " + lib_name + " time:").ljust(w) + "{:.6f} seconds".format(sum(t_lib) / args.runs))
        print(("Average " + my_name + " time:").ljust(w) + "{:.6f} seconds".format(sum(t_mine) / args.runs))
    return all_ok


def main(argv=None) -> int:
    from mpi4py import MPI
    from mpi_wrapper import Communicator

    comm = Communicator(MPI.COMM_WORLD)
    nprocs, rank = comm.Get_size(), comm.Get_rank()
    args = build_parser().parse_args(argv)
    if args.device:
        import torch
        torch.cuda.set_device(rank % max(1, torch.cuda.device_count()))
    A = _Arrays(args.device, rank)
    case = args.test_case
    ok = True

    if case == "allreduce":
        r, rr = A.randint(100, args.count), A.empty(args.count)
        print("Rank " + str(rank) + ": " + A.show(r))
        comm.Barrier()
        comm.Allreduce(r, rr, op=MPI.MIN)
        if rank == 0:
            print("Allreduce: " + A.show(rr))
        # the reference's if/elif layout also falls into the default branch for this case
        print(f"This is rank {rank}.")
    elif case == "myallreduce":
        ok = _compare(comm, MPI, args, A,
                      lambda: (A.randint(100, args.count), A.empty(args.count), A.empty(args.count)),
                      lambda s, d: comm.Allreduce(s, d, op=MPI.MIN),
                      lambda s, d: comm.myAllreduce(s, d, op=MPI.MIN),
                      "MPI.Allreduce", "myAllreduce")
    elif case == "allgather":
        r, rr = A.randint(100, 2), A.empty(2 * nprocs)
        print("Rank " + str(rank) + ": " + A.show(r))
        comm.Barrier()
        comm.Allgather(r, rr)
        if rank == 0:
            print("Allgather: " + A.show(rr))
    elif case == "reduce_scatter":
        r, rr = A.randint(100, 2 * nprocs), A.empty(2)
        print("Rank " + str(rank) + ": " + A.show(r))
        comm.Barrier()
        comm.Reduce_scatter(r, rr, op=MPI.MIN)
        print("Rank " + str(rank) + " After Reduce_scatter: " + A.show(rr))
    elif case == "split":
        # groups by rank % 4: (0,4) (1,5) (2,6) (3,7) on 8 ranks (figs/split.jpeg)
        r, rr = A.randint(100, 10), A.empty(10)
        print("Rank " + str(rank) + ": " + A.show(r))
        group = comm.Split(key=rank, color=rank % 4)
        group.Barrier()
        group.Allreduce(r, rr, op=MPI.MIN)
        print("Rank " + str(rank) + " After split and Allreduce: " + A.show(rr))
    elif case == "alltoall":
        send, recv = A.array([rank * 100 + i for i in range(nprocs)]), A.empty(nprocs)
        print("Rank " + str(rank) + " sending: " + A.show(send))
        comm.Barrier()
        comm.Alltoall(send, recv)
        print("Rank " + str(rank) + " received: " + A.show(recv))
    elif case == "myalltoall":
        ok = _compare(comm, MPI, args, A,
                      lambda: (A.array([rank * 100 + i for i in range(nprocs)]), A.empty(nprocs), A.empty(nprocs)),
                      lambda s, d: comm.Alltoall(s, d),
                      lambda s, d: comm.myAlltoall(s, d),
                      "MPI.Alltoall", "myAlltoall")
    else:
        print(f"This is rank {rank}.")
    return 0 if ok else 1
