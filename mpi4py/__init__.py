"""In-repo stand-in for the ``mpi4py`` package.

The reference imports ``from mpi4py import MPI`` everywhere (mpi_wrapper/comm.py:1,
model/func_impl.py:2, mpi-test.py:1, every MPI test at line 2), but neither mpi4py nor an MPI
runtime exists on the target image and nothing can be installed there. This package makes that
import resolve to this framework's own communicator (``collective_communication_mpi_b200.mpi``):
device kernels over NVLink when each rank has a GPU, shared memory otherwise.

If a genuine mpi4py is installed somewhere else on ``sys.path`` it wins (set
``CCMPI_FORCE_SHIM=1`` to keep the in-repo one).
"""
import importlib.machinery
import importlib.util
import os
import sys
from pathlib import Path

__version__ = "3.1.5+ccmpi_b200"


def _real_spec():
    if os.environ.get("CCMPI_FORCE_SHIM", "0") == "1":
        return None
    here = Path(__file__).resolve().parent.parent
    paths = []
    for p in sys.path:
        try:
            if p and Path(p).resolve() != here and Path(p or ".").resolve() != here:
                paths.append(p)
        except OSError:
            continue
    try:
        spec = importlib.machinery.PathFinder.find_spec("mpi4py", paths)
    except Exception:  # noqa: BLE001
        return None
    if spec is None or spec.origin is None or Path(spec.origin).resolve() == Path(__file__).resolve():
        return None
    return spec


_spec = _real_spec()
if _spec is not None:
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules[__name__] = _mod
    _spec.loader.exec_module(_mod)
else:
    def get_include():
        return str(Path(__file__).resolve().parent.parent / "include")

    def get_config():
        return {"backend": "ccmpi_b200"}
