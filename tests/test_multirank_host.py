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


def test_collectives_p2p_split_and_accounting_with_cpu_torch_tensors():
    _ok(run_ranks(4, ["tests/workers/w_collectives.py", "--torch-cpu"]), "COLLECTIVES OK ranks=4 cuda=False torch-cpu")


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


# ---- tensor-parallel layers, model and training step on CPU tensors ----------------------------
# Without a GPU the layers use the CPU statement of the fused ops (ops/host_ops.py) over the
# shared-memory transport, so the same worker that checks them on B200s checks their Python logic
# -- autograd wiring, shard bookkeeping, gradient buckets, SGD -- here, against a dense model.
@pytest.mark.parametrize("mp,dp", [(2, 2), (4, 2)])
def test_tp_layers_and_train_step_match_dense_model_on_cpu(mp, dp):
    _ok(run_ranks(mp * dp, ["tests/workers/w_layers.py", str(mp), str(dp)], timeout=240), f"LAYERS OK mp={mp} dp={dp}")


def test_train_step_with_sgd_fused_into_the_weight_gradient_on_cpu():
    """dp = 1: every dW GEMM applies its own SGD step; dx must have read W before that."""
    p = run_ranks(2, ["tests/workers/w_layers.py", "2", "1"], timeout=240, env={"CCMPI_FUSED_SGD": "1"})
    _ok(p, "LAYERS OK mp=2 dp=1")
    assert "fused_sgd=1" in p.stdout
    # with data parallelism the switch is ignored (the gradient has to be reduced first)
    p = run_ranks(2, ["tests/workers/w_layers.py", "1", "2"], timeout=240, env={"CCMPI_FUSED_SGD": "1"})
    _ok(p, "LAYERS OK mp=1 dp=2")
    assert "fused_sgd=0" in p.stdout


def test_train_step_with_the_fused_elementwise_ops_on_cpu():
    _ok(run_ranks(2, ["tests/workers/w_layers.py", "2", "1"], timeout=240, env={"CCMPI_FUSED_ELEMENTWISE": "1"}),
        "LAYERS OK mp=2 dp=1")


@pytest.mark.parametrize("nranks", [4])
def test_column_parallel_layer_with_gathered_output_on_cpu(nranks):
    _ok(run_ranks(nranks, ["tests/workers/w_gather_output.py"], timeout=240), f"GATHER OUTPUT OK ranks={nranks}")


TORCH_CPU_CODE = r'''
import torch
from mpi4py import MPI
w = MPI.COMM_WORLD
r, n = w.Get_rank(), w.Get_size()
for dt in (torch.float32, torch.bfloat16, torch.int64):
    x = (torch.arange(2 * 3 * 4).reshape(2, 3, 4) + 100 * r).to(dt)
    g = w.gather_last_axis(x)                                        # func_impl.py:76-91 on CPU tensors
    assert g.dtype == dt and g.shape == (2, 3, 4 * n)
    want = torch.cat([(torch.arange(24).reshape(2, 3, 4) + 100 * q).to(dt) for q in range(n)], dim=-1)
    assert torch.equal(g, want)
    p = (torch.arange(2 * 3 * 2 * n).reshape(2, 3, 2 * n) + r).to(dt)
    s = w.reduce_scatter_last_axis(p)                                # func_impl.py:150-187
    total = sum((torch.arange(2 * 3 * 2 * n).reshape(2, 3, 2 * n) + q).float() for q in range(n))
    assert s.dtype == dt and torch.equal(s.float(), total[..., 2 * r:2 * r + 2].to(dt).float())
w.Barrier()
if r == 0: print("TORCH CPU OK", n)
'''


def test_last_axis_collectives_accept_cpu_torch_tensors():
    _ok(run_ranks(4, ["-c", TORCH_CPU_CODE], timeout=120), "TORCH CPU OK 4")


GATHER_SCATTER_CODE = r'''
import numpy as np
from mpi4py import MPI
w = MPI.COMM_WORLD; r, n = w.Get_rank(), w.Get_size()
for root in (0, n - 1):
    out = np.empty(3 * n, dtype=np.float64) if r == root else None
    w.Gather(np.full(3, r + 0.5), out, root=root)
    if r == root:
        assert out.tolist() == [q + 0.5 for q in range(n) for _ in range(3)]
    src = (np.arange(2 * n, dtype=np.int32) + 100 * root) if r == root else None
    got = np.empty(2, dtype=np.int32)
    w.Scatter(src, got, root=root)
    assert got.tolist() == [2 * r + 100 * root, 2 * r + 1 + 100 * root], (r, got)
nxt, prv = (r + 1) % n, (r - 1) % n
got = w.sendrecv({"from": r, "pad": "x" * (r * 1000)}, dest=nxt, source=prv)
assert got["from"] == prv and len(got["pad"]) == prv * 1000
w.Barrier()
if r == 0: print("GS OK", n)
'''


def test_gather_scatter_and_object_sendrecv():
    """mpi4py members beyond what the reference touches, for users switching over."""
    _ok(run_ranks(4, ["-c", GATHER_SCATTER_CODE], timeout=120), "GS OK 4")


def test_training_example_runs_and_learns_on_cpu():
    import sys
    from conftest import REPO
    import subprocess
    p = subprocess.run([sys.executable, "-m", "collective_communication_mpi_b200.launch", "-n", "2", sys.executable,
                        "examples/train_block.py", "--mp", "2", "--steps", "6"], cwd=REPO, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "TRAIN OK" in p.stdout, p.stdout[-1000:]


@pytest.mark.parametrize("mp,dp,env", [(2, 2, None), (2, 1, {"CCMPI_FUSED_SGD": "1", "CCMPI_FUSED_ELEMENTWISE": "1"})])
def test_sequence_parallel_block_matches_dense_model_on_cpu(mp, dp, env):
    """Megatron-SP layout (parallel/sp_layers.py): token-sharded activations outside the GEMM pairs."""
    _ok(run_ranks(mp * dp, ["tests/workers/w_layers.py", str(mp), str(dp), "--sp"], timeout=240, env=env),
        f"LAYERS OK mp={mp} dp={dp} sp=1")


@pytest.mark.parametrize("mp,dp", [(2, 2), (4, 1)])
def test_megatron_layout_block_matches_dense_model_on_cpu(mp, dp):
    """Classic Megatron layout (parallel/megatron_layers.py): replicated activations, GEMM -> all-reduce."""
    _ok(run_ranks(mp * dp, ["tests/workers/w_layers.py", str(mp), str(dp), "--megatron"], timeout=240),
        f"LAYERS OK mp={mp} dp={dp} sp=0 megatron=1")


def test_sibling_groups_that_split_again_get_distinct_ids_and_do_not_disturb_each_other():
    """Two halves of the world split at the same time: ids come from one job-wide counter."""
    _ok(run_ranks(8, ["tests/workers/w_nested_split.py"], timeout=120), "NESTED SPLIT OK")


@pytest.mark.parametrize("flags", [[], ["--sp"], ["--megatron"]], ids=["feature-sharded", "sequence-parallel", "megatron"])
def test_blocks_can_be_given_another_token_count(flags):
    """``set_tokens``: fused ops closed and rebuilt, weights kept, results still those of the dense model."""
    sp, meg = int("--sp" in flags), int("--megatron" in flags)
    _ok(run_ranks(2, ["tests/workers/w_set_tokens.py", "2"] + flags, timeout=240), f"SET TOKENS OK mp=2 sp={sp} megatron={meg}")


@pytest.mark.parametrize("n", [3, 8])
def test_status_probe_and_any_source(n):
    """mpi4py's ``Status`` / ``Probe`` / ``Iprobe`` / ``ANY_SOURCE`` on the host plane (master/worker pattern)."""
    _ok(run_ranks(n, ["tests/workers/w_probe_status.py"], timeout=120), f"PROBE STATUS OK ranks={n}")


@pytest.mark.parametrize("n,flags", [(4, []), (8, []), (3, ["--torch"])])
def test_v_collectives_with_ragged_blocks(n, flags):
    """``Allgatherv`` / ``Gatherv`` / ``Scatterv`` / ``Alltoallv`` / ``Reduce_scatter(recvcounts=...)``."""
    _ok(run_ranks(n, ["tests/workers/w_vcollectives.py"] + flags, timeout=120), f"V COLLECTIVES OK ranks={n} torch={int(bool(flags))}")


@pytest.mark.parametrize("mp,dp,depth", [(2, 1, 3), (2, 2, 2), (4, 1, 2)])
def test_block_stack_with_one_shared_scatter_op_matches_dense_model(mp, dp, depth):
    """``TransformerFCStack``: several blocks, every reduce-scatter of the stack through ONE op object."""
    _ok(run_ranks(mp * dp, ["tests/workers/w_stack.py", str(mp), str(dp), str(depth)], timeout=300), f"STACK OK mp={mp} dp={dp} depth={depth}")


def test_examples_stack_and_task_farm():
    import os
    import subprocess
    import sys
    from conftest import REPO
    env = dict(os.environ, PYTHONPATH=REPO)
    base = [sys.executable, "-m", "collective_communication_mpi_b200.launch", "--transport", "host"]
    p = subprocess.run(base + ["-n", "4", sys.executable, "examples/train_block.py", "--mp", "2", "--steps", "5", "--depth", "2"],
                       cwd=REPO, capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0 and "TRAIN OK" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]
    p = subprocess.run(base + ["-n", "5", sys.executable, "examples/master_worker.py", "--items", "30"],
                       cwd=REPO, capture_output=True, text=True, timeout=120, env=env)
    assert p.returncode == 0 and "FARM OK" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]
