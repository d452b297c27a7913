"""Test-suite configuration: markers and shared helpers.

``gpu``  -- needs a CUDA device; the CPU-only CI runs ``-m "not gpu"``.
``mpi``  -- needs several ranks; runs only when pytest itself is started under the launcher with
            ``--with-mpi`` (the multi-rank suites are otherwise driven from single-process tests
            that spawn the launcher, see tests/test_multirank_host.py).
"""
import os
import subprocess
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (B200)")
    config.addinivalue_line("markers", "mpi: test needs several ranks; runs only with --with-mpi")


def run_ranks(nprocs, args, transport="host", timeout=300, env=None):
    """Run ``python <args>`` as ``nprocs`` ranks under the launcher; returns CompletedProcess."""
    cmd = [sys.executable, "-m", "collective_communication_mpi_b200.launch", "-n", str(nprocs),
           "--transport", transport, "--timeout", str(timeout), "--", sys.executable] + list(args)
    e = dict(os.environ)
    e["PYTHONPATH"] = REPO + (os.pathsep + e["PYTHONPATH"] if e.get("PYTHONPATH") else "")
    e.pop("RANK", None); e.pop("WORLD_SIZE", None)
    if env:
        e.update(env)
    return subprocess.run(cmd, cwd=REPO, env=e, capture_output=True, text=True, timeout=timeout + 60)


@pytest.fixture
def ranks():
    return run_ranks


def gpu_count() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:  # noqa: BLE001
        return 0
