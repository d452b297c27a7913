"""The reference's own test files, byte-for-byte, against this framework.

``scripts/install_reference.sh`` puts the unmodified reference under ``baseline/_ref`` (falling
back to ``/root/reference`` when that mount exists). Its four test files are run from there, with
THIS repository first on ``sys.path`` so that ``mpi4py``, ``model.func_impl``, ``mpi_wrapper`` and
``data.data_parallel_preprocess`` resolve to our implementations, at the rank counts the
reference's README prescribes (README.md:163,189,200,212).
"""
import os

import pytest

from conftest import REPO, run_ranks

CANDIDATES = [os.path.join(REPO, "baseline", "_ref", "tests"), "/root/reference/tests"]
REF_TESTS = next((d for d in CANDIDATES if os.path.isfile(os.path.join(d, "test_get_info.py"))), None)

pytestmark = pytest.mark.skipif(REF_TESTS is None, reason="reference tests not present (run scripts/install_reference.sh)")

PYTEST = ["-m", "pytest", "-l", "-q", "-p", "collective_communication_mpi_b200.pytest_plugin", "-p", "no:cacheprovider",
          "--rootdir", "/tmp"]


def _run(nranks, fname, transport, expect, env=None):
    p = run_ranks(nranks, PYTEST + (["--with-mpi"] if nranks > 1 else []) + [os.path.join(REF_TESTS, fname)],
                  transport=transport, env={"CCMPI_FORCE_SHIM": "1", **(env or {})})
    assert p.returncode == 0, f"{fname} on {nranks} ranks ({transport})\n{p.stdout[-4000:]}\n{p.stderr[-2000:]}"
    assert expect in p.stdout, p.stdout[-2000:]


def test_reference_data_split_file():
    _run(1, "test_data_split.py", "host", "4 passed")


def test_reference_get_info_file_on_8_ranks():
    _run(8, "test_get_info.py", "host", "2 passed")


def test_reference_forward_file_on_4_ranks():
    _run(4, "test_transformer_forward.py", "host", "2 passed")


def test_reference_forward_file_on_8_ranks_with_empty_shards():
    _run(8, "test_transformer_forward.py", "host", "2 passed")


def test_reference_backward_file_on_4_ranks():
    _run(4, "test_transformer_backward.py", "host", "2 passed")


# ---- the same files on the DEVICE plane: every collective they make is a CUDA kernel over NVLink ----
# (int64 SUM all-reduce over mp_comm / dp_comm, fp64 last-axis gather and reduce-scatter through the
# strided all-gather / reduce-scatter kernels; NumPy arrays are staged through the symmetric heap).
# The tests' arrays are tiny; by default such host payloads would stay on the shared-memory plane
# (mpi/core.py: _DeviceSide.agree), so the route is switched off here to keep the kernels under test.
KERNELS = {"CCMPI_HOST_ROUTE_MAX": "0"}
def _need_gpus(n):
    from conftest import gpu_count
    if gpu_count() < n:
        pytest.skip(f"needs {n} GPUs")


@pytest.mark.gpu
def test_reference_get_info_file_on_8_gpus_device_plane():
    _need_gpus(8)
    _run(8, "test_get_info.py", "device", "2 passed", KERNELS)


@pytest.mark.gpu
def test_reference_forward_file_on_4_gpus_device_plane():
    _need_gpus(4)
    _run(4, "test_transformer_forward.py", "device", "2 passed", KERNELS)


@pytest.mark.gpu
def test_reference_backward_file_on_4_gpus_device_plane():
    _need_gpus(4)
    _run(4, "test_transformer_backward.py", "device", "2 passed", KERNELS)


@pytest.mark.gpu
def test_reference_forward_file_on_8_gpus_device_plane_with_empty_shards():
    _need_gpus(8)
    _run(8, "test_transformer_forward.py", "device", "2 passed", KERNELS)


@pytest.mark.gpu
def test_reference_files_on_the_device_plane_with_default_routing():
    """As a user runs them: device transport up, tiny NumPy payloads routed to the host plane."""
    _need_gpus(4)
    _run(4, "test_transformer_forward.py", "device", "2 passed")
    _run(4, "test_transformer_backward.py", "device", "2 passed")


@pytest.mark.parametrize("n", [4, 8])
def test_differential_against_the_unmodified_reference_modules(n):
    """Reference ``Communicator`` / ``func_impl`` / ``split_data`` loaded from baseline/_ref and run side by side
    with ours on the same communicators: equal results bit for bit, equal byte accounting (w_differential.py)."""
    p = run_ranks(n, ["tests/workers/w_differential.py"], transport="host", timeout=240, env={"CCMPI_FORCE_SHIM": "1"})
    assert p.returncode == 0, f"{p.stdout[-3000:]}\n{p.stderr[-3000:]}"
    assert f"DIFFERENTIAL OK ranks={n}" in p.stdout


CASES = [None, "allreduce", "myallreduce", "allgather", "reduce_scatter", "split", "alltoall", "myalltoall"]


@pytest.mark.parametrize("case", CASES, ids=[c or "default" for c in CASES])
def test_cli_prints_what_the_reference_cli_prints(case):
    """``mpi-test.py`` of the reference and ours, 8 ranks each: the same lines once numbers are blanked (payloads are
    random, times differ) and rank-prefixed lines are sorted (ranks print concurrently)."""
    import re
    ref_cli = os.path.join(os.path.dirname(REF_TESTS), "mpi-test.py")
    if not os.path.isfile(ref_cli):
        pytest.skip("reference mpi-test.py not present")
    args = ["--test_case", case] if case else []

    def shape(script):
        p = run_ranks(8, [script] + args, transport="host", timeout=240, env={"CCMPI_FORCE_SHIM": "1"})
        assert p.returncode == 0, f"{script}\n{p.stdout[-3000:]}\n{p.stderr[-3000:]}"
        lines = [re.sub(r"-?\d+(\.\d+)?(e[-+]?\d+)?", "#", ln).strip() for ln in p.stdout.splitlines() if ln.strip()]
        lines = [re.sub(r"\[\s+", "[", re.sub(r"\s+", " ", ln)) for ln in lines]      # numpy pads short numbers
        return sorted(lines)

    ours, theirs = shape("mpi-test.py"), shape(ref_cli)
    # ours may add lines (e.g. the summary of --runs); every line the reference prints must be there, as often
    from collections import Counter
    missing = Counter(theirs) - Counter(ours)
    assert not missing, f"lines of the reference's output that ours lacks: {dict(missing)}\nours: {ours}"
