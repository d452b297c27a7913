"""``mpirun -n N`` replacement.

    python -m collective_communication_mpi_b200.launch -n 8 python mpi-test.py --test_case myallreduce
    python -m collective_communication_mpi_b200.launch -n 4 python -m pytest -l -v --with-mpi tests/...

Starts N copies of the command on this node, one per rank (and per GPU when there are GPUs),
with RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT / CCMPI_JOB exported -- the same
variables ``torchrun`` sets, so either launcher works. Mirrors what the reference gets from
``mpirun -n 8 python ...`` (README.md:52,130,136,189,200,212).

Behaviour that matters for tests: every rank's output is passed through (optionally tagged
with ``[rank]``), the exit code is the first non-zero one, and when any rank dies the rest are
terminated instead of being left waiting in a collective.
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import threading
import time
import uuid


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(nprocs: int, cmd, *, tag_output: bool = False, env_extra=None, timeout: float | None = None,
           capture: bool = False):
    """Run ``cmd`` (list of str) as ``nprocs`` ranks. Returns ``(exit_code, outputs)`` where
    ``outputs`` is a per-rank list of captured text when ``capture`` is set."""
    job = uuid.uuid4().hex[:12]
    port = _free_port()
    procs = []
    outputs = [[] for _ in range(nprocs)]
    threads = []
    out_lock = threading.Lock()
    repo_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for r in range(nprocs):
        env = dict(os.environ)
        env.update({
            "RANK": str(r), "WORLD_SIZE": str(nprocs), "LOCAL_RANK": str(r), "LOCAL_WORLD_SIZE": str(nprocs),
            "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "CCMPI_JOB": job,
            "PYTHONUNBUFFERED": "1",
        })
        env["PYTHONPATH"] = repo_root + (os.pathsep + env["PYTHONPATH"] if env.get("PYTHONPATH") else "")
        if nprocs > 1:
            # like torchrun: N ranks that each start one compute thread per core thrash the node;
            # give every rank its share unless the user has said otherwise
            env.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // nprocs)))
        if env_extra:
            env.update(env_extra)
        # every rank's stdout / stderr comes through a pipe and is forwarded one whole line at a time
        # under a lock (what mpirun's I/O forwarding does): ranks that print at the same moment cannot
        # splice their lines into each other, which they do when N children share the terminal directly
        p = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                             start_new_session=True, errors="replace")
        procs.append(p)

        def pump(src, dst, rank=r, keep=True):
            for line in src:
                if keep:
                    outputs[rank].append(line)
                if not capture or tag_output:
                    with out_lock:
                        dst.write(f"[{rank}] {line}" if tag_output else line)
                        dst.flush()
        for src, dst, keep in ((p.stdout, sys.stdout, True), (p.stderr, sys.stderr, capture)):
            t = threading.Thread(target=pump, args=(src, dst), kwargs={"keep": keep}, daemon=True)
            t.start()
            threads.append(t)

    def kill_all():
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except (ProcessLookupError, PermissionError):
                    pass
        t_end = time.time() + 5
        for p in procs:
            while p.poll() is None and time.time() < t_end:
                time.sleep(0.05)
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except (ProcessLookupError, PermissionError):
                    pass

    rc = 0
    t0 = time.time()
    try:
        live = set(range(nprocs))
        while live:
            for r in list(live):
                code = procs[r].poll()
                if code is not None:
                    live.discard(r)
                    if code != 0 and rc == 0:
                        rc = code
                        sys.stderr.write(f"[launch] rank {r} exited with code {code}; stopping the job\n")
                        kill_all()
                        live.clear()
                        break
            if timeout is not None and time.time() - t0 > timeout and live:
                sys.stderr.write(f"[launch] timeout after {timeout:.0f} s; stopping the job\n")
                rc = 124
                kill_all()
                break
            time.sleep(0.02)
    except KeyboardInterrupt:
        rc = 130
        kill_all()
    for t in threads:
        t.join(timeout=2)
    # a rank that died may have left the job's shared-memory segment behind
    try:
        os.unlink(f"/dev/shm/ccmpi-{job}")
    except OSError:
        pass
    return rc, ["".join(o) for o in outputs]


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m collective_communication_mpi_b200.launch",
                                 description="Start N ranks of a command on this node (mpirun -n N).")
    ap.add_argument("-n", "-np", "--nprocs", type=int, required=True, dest="nprocs")
    ap.add_argument("--tag-output", action="store_true", help="prefix every line with [rank]")
    ap.add_argument("--timeout", type=float, default=None, help="kill the job after this many seconds")
    ap.add_argument("--transport", choices=["auto", "host", "device"], default=None,
                    help="force the data plane (default: CCMPI_TRANSPORT or auto)")
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    args = ap.parse_args(argv)
    cmd = args.cmd[1:] if args.cmd[:1] == ["--"] else args.cmd
    if not cmd:
        ap.error("no command given")
    extra = {"CCMPI_TRANSPORT": args.transport} if args.transport else None
    rc, _ = launch(args.nprocs, cmd, tag_output=args.tag_output, env_extra=extra, timeout=args.timeout)
    sys.exit(rc)


if __name__ == "__main__":
    main()
