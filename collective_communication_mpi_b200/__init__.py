"""collective_communication_mpi_b200 -- a B200-native 2-D (data x tensor-model parallel)
communication framework with the capabilities and public API of
anaykulkarni/collective-communication-mpi.

Layout:

* ``runtime/``   process bootstrap, symmetric heap, device communicators, launcher
* ``mpi/``       the ``mpi4py.MPI``-compatible front end (host and device transports)
* ``ops/``       device collectives, tcgen05 GEMM, fused GEMM+collective kernels
* ``parallel/``  2-D layout (``get_info``), TP FC layers, DP gradient reduction
* ``models/``    transformer FC stack / block, MLP stack
* ``utils/``     NumPy oracle, timing, clocks

The reference's own import paths (``mpi_wrapper``, ``model.func_impl``,
``data.data_parallel_preprocess``, ``mpi4py``) are provided as top-level packages of this
repository and are thin shims over this package.
"""
__version__ = "0.1.0"
