from .comm import Communicator  # noqa: F401  (same export as the reference's mpi_wrapper/__init__.py:1)
