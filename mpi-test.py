"""Demo / micro-benchmark CLI with the reference's name and flags (reference mpi-test.py):

    python -m collective_communication_mpi_b200.launch -n 8 python mpi-test.py --test_case myallreduce

The cases are implemented in ``collective_communication_mpi_b200/cli.py``.
"""
import sys

from collective_communication_mpi_b200.cli import main

if __name__ == "__main__":
    sys.exit(main())
