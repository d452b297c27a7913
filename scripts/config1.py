"""BASELINE config 1 as data: the reference's own micro-benchmark (``mpi-test.py --test_case myallreduce`` /
``myalltoall``, 100 runs of 100 x int64 / n x int64, mpi-test.py:40-98,178-239) on this box's CPUs, next to
this framework's ``mpi-test.py`` on the host plane and, when the box has a GPU per rank, on the device plane.

    python scripts/config1.py [-n 8] [--out profiles/r2_config1_<where>.jsonl]

Arms (one JSON line each per test case; times are the averages the programs print, in microseconds):
  reference/host-shim   baseline/_ref/mpi-test.py, unmodified, over the in-repo mpi4py stand-in (no MPI on the image)
  ours/host             this repository's mpi-test.py, NumPy arrays, shared-memory transport
  ours/device-staged    the same with the device transport: NumPy arrays staged through the symmetric heap
  ours/device           --device: arrays live on the GPUs, the timed region is kernel + stream sync
"""
import argparse
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(n, script, case, transport, extra=(), cwd=ROOT):
    cmd = [sys.executable, "-m", "collective_communication_mpi_b200.launch", "-n", str(n), "--transport", transport, "--",
           sys.executable, script, "--test_case", case, *extra]
    env = dict(os.environ, PYTHONPATH=ROOT, CCMPI_FORCE_SHIM="1")
    p = subprocess.run(cmd, cwd=cwd, env=env, capture_output=True, text=True, timeout=600)
    avg = dict(re.findall(r"Average (\S+) time:\s+([0-9.eE+-]+) seconds", p.stdout))
    ok = "All runs produced correct results." in p.stdout
    return p.returncode, ok, {k: float(v) * 1e6 for k, v in avg.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-n", type=int, default=8)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    try:
        import torch
        gpus = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:  # noqa: BLE001
        gpus = 0
    ref_script = os.path.join(ROOT, "baseline", "_ref", "mpi-test.py")
    arms = [("reference/host-shim", ref_script, "host", ()), ("ours/host", "mpi-test.py", "host", ())]
    if gpus >= args.n:
        arms += [("ours/device-staged", "mpi-test.py", "device", ()), ("ours/device", "mpi-test.py", "device", ("--device",))]
    rows = []
    for case in ("myallreduce", "myalltoall"):
        for arm, script, transport, extra in arms:
            if not os.path.isfile(os.path.join(ROOT, script)) and not os.path.isfile(script):
                continue
            rc, ok, avg = run(args.n, script, case, transport, extra)
            lib = "MPI.Allreduce" if case == "myallreduce" else "MPI.Alltoall"
            mine = "myAllreduce" if case == "myallreduce" else "myAlltoall"
            rows.append({"config": "1", "case": case, "arm": arm, "ranks": args.n, "runs": 100, "rc": rc, "all_correct": ok,
                         "library_call_us": round(avg.get(lib, float("nan")), 2), "my_call_us": round(avg.get(mine, float("nan")), 2),
                         "cpus": os.cpu_count(), "gpus": gpus})
            print(json.dumps(rows[-1]), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
