"""Multi-rank behaviour on the CPU shared-memory plane (runs anywhere, no GPU).

Each test starts N ranks with the launcher (the ``mpirun -n N`` replacement) and lets the worker
scripts in tests/workers/ assert on every rank; the launcher exits non-zero if any rank fails.
"""
import pytest

from conftest import run_ranks


def _ok(proc, token):
    assert proc.returncode == 0, f"rc={proc.returncode}\n--- stdout\n{proc.stdout[-3000:]}\n--- stderr\n{proc.stderr[-3000:]}"
    assert token in proc.stdout, proc.stdout[-2000:]


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_collectives_p2p_split_and_accounting(nranks):
    _ok(run_ranks(nranks, ["tests/workers/w_collectives.py"]), f"COLLECTIVES OK ranks={nranks}")


@pytest.mark.parametrize("nranks", [4, 8])
def test_grid_and_tp_communication_steps(nranks):
    _ok(run_ranks(nranks, ["tests/workers/w_tp_steps.py"]), f"TP STEPS OK ranks={nranks}")


@pytest.mark.parametrize("case,expect", [
    ("allreduce", "Allreduce: ["), ("allgather", "Allgather: ["), ("reduce_scatter", "After Reduce_scatter"),
    ("split", "After split and Allreduce"), ("alltoall", "Rank 7 received: [  7 107 207 307 407 507 607 707]"),
    ("myallreduce", "All runs produced correct results."), ("myalltoall", "All runs produced correct results."),
    ("", "This is rank 3."),
])
def test_cli_cases_on_eight_ranks(case, expect):
    args = ["mpi-test.py"] + (["--test_case", case, "--runs", "10"] if case else [])
    p = run_ranks(8, args)
    _ok(p, expect)
    if case in ("myallreduce", "myalltoall"):
        assert "Average my" in p.stdout and "Summary over 10 runs:" in p.stdout


def test_a_failing_rank_stops_the_job_instead_of_hanging():
    code = ("from mpi4py import MPI\nimport sys\nc = MPI.COMM_WORLD\n"
            "if c.Get_rank() == 1: sys.exit(3)\nc.Barrier()\n")
    p = run_ranks(4, ["-c", code], timeout=60, env={"CCMPI_TIMEOUT": "20"})
    assert p.returncode != 0


def test_pytest_itself_runs_under_the_launcher_with_the_mpi_flag(tmp_path):
    t = tmp_path / "test_ranked.py"
    t.write_text(
        "import pytest\nfrom mpi4py import MPI\n\n"
        "@pytest.mark.mpi\ndef test_sum():\n"
        "    c = MPI.COMM_WORLD\n    assert c.allreduce(1) == c.Get_size() == 3\n")
    p = run_ranks(3, ["-m", "pytest", "-q", "--with-mpi", "-p", "collective_communication_mpi_b200.pytest_plugin",
                      "--rootdir", str(tmp_path), str(t)])
    _ok(p, "1 passed")
    p = run_ranks(1, ["-m", "pytest", "-q", "-rs", "-p", "collective_communication_mpi_b200.pytest_plugin",
                      "--rootdir", str(tmp_path), str(t)])
    assert "1 skipped" in p.stdout, p.stdout
