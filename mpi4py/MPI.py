"""``mpi4py.MPI`` surface backed by collective_communication_mpi_b200 (see package docstring).

What the reference and its tests touch: ``COMM_WORLD``, ``Comm``,
``SUM/MIN/MAX``, ``Wtime``, ``Request.Waitall`` and the ``Comm`` methods. The world communicator
is created on first access so that merely importing this module starts nothing.
"""
from collective_communication_mpi_b200.mpi.core import (  # noqa: F401
    ANY_SOURCE, ANY_TAG, BYTE, CHAR, DOUBLE, FLOAT, IN_PLACE, INT, INT32_T, INT64_T, LAND, LONG, LOR, MAX, MIN,
    PROC_NULL, PROD, SUM, UNDEFINED, Comm, Datatype, Finalize, Get_processor_name, Is_finalized, Is_initialized,
    Op, Request, Status, Wtime, comm_world,
)

Intracomm = Comm
COMM_NULL = None


def Get_version():
    return (3, 1)


def __getattr__(name):
    if name == "COMM_WORLD":
        return comm_world()
    if name == "COMM_SELF":
        w = comm_world()
        return w.Split(color=w.Get_rank(), key=0)
    raise AttributeError(f"module 'mpi4py.MPI' (ccmpi_b200 front end) has no attribute {name!r}")
