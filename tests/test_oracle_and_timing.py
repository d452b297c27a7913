"""Oracles, byte accounting formulas and timing helpers (CPU)."""
import numpy as np
import pytest

from conftest import run_ranks
from collective_communication_mpi_b200.utils import oracle as O
from collective_communication_mpi_b200.utils.timing import OpStats, bus_bandwidth_gbs, time_host


def test_accounting_oracles_of_the_reference_at_eight_ranks():
    # concrete values for n = 8
    assert O.bytes_accounted("Allreduce", 800, 800, 8) == 11200
    assert O.bytes_accounted("Allgather", 16, 128, 8) == 1008
    assert O.bytes_accounted("Reduce_scatter", 128, 16, 8) == 1008
    assert O.bytes_accounted("Alltoall", 64, 64, 8) == 112
    assert O.bytes_accounted("myAllreduce", 800, 800, 8, rank=0) == 11200
    assert O.bytes_accounted("myAllreduce", 800, 800, 8, rank=3) == 1600
    assert O.bytes_accounted("myAlltoall", 64, 64, 8) == 112


def test_collective_oracles_on_small_cases():
    parts = [np.array([1, 5, 2, 8]), np.array([4, 0, 9, 3])]
    assert O.allreduce(parts, "min").tolist() == [1, 0, 2, 3]
    assert O.allgather(parts).tolist() == [1, 5, 2, 8, 4, 0, 9, 3]
    assert [p.tolist() for p in O.reduce_scatter(parts)] == [[5, 5], [11, 11]]
    assert [p.tolist() for p in O.alltoall(parts)] == [[1, 5, 4, 0], [2, 8, 9, 3]]
    x = [np.ones((1, 2, 3)), 2 * np.ones((1, 2, 3))]
    assert O.gather_last_axis(x).shape == (1, 2, 6)
    assert O.reduce_scatter_last_axis([np.arange(4.0).reshape(1, 1, 4)] * 2)[1].tolist() == [[[4.0, 6.0]]]


def test_bus_bandwidth_convention_and_stats():
    assert bus_bandwidth_gbs("allreduce", 10 ** 9, 1.0, 8) == pytest.approx(1.75)
    assert bus_bandwidth_gbs("allgather", 10 ** 9, 1.0, 8) == pytest.approx(0.875)
    st = OpStats()
    st.add("Allreduce", 1000, 1e-6)
    st.add("Allreduce", 1000, 1e-6)
    r = st.report()["Allreduce"]
    assert r["calls"] == 2 and r["bytes"] == 2000 and r["GBps"] == pytest.approx(1.0)
    assert time_host(lambda: None, iters=5, warmup=1) >= 0.0


@pytest.mark.parametrize("nranks", [2, 4])
def test_transports_and_tp_layers_against_the_oracle(nranks):
    p = run_ranks(nranks, ["tests/workers/w_oracle.py"])
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert f"ORACLE OK ranks={nranks}" in p.stdout


def test_committed_result_tables_are_what_the_raw_measurement_files_say():
    """profiles/R2_RESULTS.md is generated from the raw r2_* files (scripts/make_profiles_readme.py); a table that
    was edited by hand, or raw files changed without regenerating, fails here."""
    import os
    import subprocess
    import sys
    from conftest import REPO
    p = subprocess.run([sys.executable, os.path.join("scripts", "make_profiles_readme.py"), "--check"], cwd=REPO,
                       capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
