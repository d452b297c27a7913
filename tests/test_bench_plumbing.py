"""bench.py contract checks that need no GPU: the reference arm runs the unmodified reference from
baseline/_ref (tiny smoke override) and prints one well-formed JSON line; the launcher exports the
torchrun-style environment."""
import json
import os
import subprocess
import sys

import pytest

from conftest import REPO, run_ranks

HAVE_REF = os.path.isfile(os.path.join(REPO, "baseline", "_ref", "model", "func_impl.py"))


def _bench(args, env_extra, nproc=1):
    env = dict(os.environ, CCMPI_BENCH_TINY="1", **env_extra)
    env.pop("RANK", None); env.pop("WORLD_SIZE", None)
    if nproc == 1:
        return subprocess.run([sys.executable, "bench.py"] + args, cwd=REPO, env=env, capture_output=True, text=True, timeout=600)
    return run_ranks(nproc, ["bench.py"] + args, env={"CCMPI_BENCH_TINY": "1", **env_extra}, timeout=600)


@pytest.mark.skipif(not HAVE_REF, reason="baseline/_ref not installed")
@pytest.mark.parametrize("nproc", [1, 2])
def test_reference_arm_emits_one_json_line(nproc):
    p = _bench(["--impl", "reference", "--gpus", str(nproc), "--steps", "2", "--warmup", "1"], {}, nproc)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == nproc
    assert d["metric"] == "tp_fc_fwd_bwd_tokens_per_s" and d["value"] > 0 and d["higher_is_better"] is True
    assert "shared-memory stand-in" in d["mpi_runtime"]
    # none of this repository's kernels or engine on the reference arm: only the host transport is mapped
    libs = d["native_libs_mapped"]      # (a sanitizer build of the host transport has another file name)
    assert len(libs) == (1 if nproc > 1 else 0) and all(l.startswith("libccmpi_host") for l in libs), libs
    assert d["gpu_launches"] == 0 and d["e2e"]["ms_per_step"] == d["ms_per_step"]


def test_reference_arm_shortens_the_run_to_its_wall_budget_and_says_so():
    if not HAVE_REF:
        pytest.skip("baseline/_ref not installed")
    p = _bench(["--impl", "reference", "--steps", "1000000", "--warmup", "5"], {"CCMPI_REF_BUDGET_S": "1"})
    assert p.returncode == 0
    d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
    assert d["impl"] == "reference" and "unavailable" not in d
    assert 1 <= d["steps"] < 1000000 and d["warmup"] >= 1 and "wall budget" in d["note"]
    assert d["value"] > 0 and d["e2e"]["ms_per_step"] == d["ms_per_step"]


def test_launcher_exports_torchrun_style_environment():
    code = ("import os\nkeys = ['RANK','WORLD_SIZE','LOCAL_RANK','MASTER_ADDR','MASTER_PORT','CCMPI_JOB']\n"
            "assert all(k in os.environ for k in keys)\nprint('ENV', os.environ['RANK'], os.environ['WORLD_SIZE'])\n")
    p = run_ranks(3, ["-c", code])
    assert p.returncode == 0 and sorted(l for l in p.stdout.splitlines() if l.startswith("ENV")) == \
        ["ENV 0 3", "ENV 1 3", "ENV 2 3"]


def test_info_command_describes_the_installation():
    import json
    import subprocess
    import sys
    p = subprocess.run([sys.executable, "-m", "collective_communication_mpi_b200.info", "--json"], cwd=REPO, capture_output=True,
                       text=True, timeout=120, env=dict(os.environ, PYTHONPATH=REPO, CCMPI_HEAP_MB="512"))
    assert p.returncode == 0, p.stderr[-2000:]
    info = json.loads(p.stdout)
    assert info["libraries"]["host"]["loads"] and info["libraries"]["device"]["loads"]
    assert info["limits"]["flag_words_for_fused_ops"] == info["limits"]["signal_pad_words"] - 2048
    assert info["environment"].get("CCMPI_HEAP_MB") == "512"
    assert isinstance(info["gpus"], list) and "device_plane" in info
