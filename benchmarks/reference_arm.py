"""Reference arm of bench.py: the UNMODIFIED reference (baseline/_ref, installed by
scripts/install_reference.sh) on the same metric and configuration.

The reference is a CPU library: ``get_info`` / ``naive_collect_forward_input`` /
``naive_collect_backward_output`` / ``naive_collect_backward_x`` / ``split_data`` / ``Communicator``
move NumPy arrays between MPI ranks; it has no GEMM and no GPU path. The only way it can produce
"TP FC fwd+bwd tokens/s" is therefore: its own communication functions, called exactly as its
README describes the layer (README.md:173-213), around NumPy matmuls on the host. That is what is
timed here (fp32 NumPy: the reference has no bf16). Nothing of this repository's models, kernels
or engine is on the path -- with one unavoidable exception that is reported in the JSON line:
``mpi4py`` and an MPI runtime do not exist on the image and cannot be installed, so
``from mpi4py import MPI`` inside the reference resolves to this repository's shared-memory
stand-in (host transport; CCMPI_TRANSPORT=host is forced so no GPU kernel runs).

``baseline/_ref`` is tracked in git (the box that runs this arm has no ``/root/reference``); if it is
missing anyway and ``/root/reference`` exists, ``scripts/install_reference.sh`` is run first. If there
is no reference at all, a single ``{"impl": "reference", "unavailable": ...}`` line is printed and the
exit code is 0, as the task statement prescribes.

Wall budget: one CPU step of this configuration takes tens of seconds (54 s on 8 cores, measured), so
the requested ``--warmup W --steps K`` may not fit the caller's time limit. The first step is timed;
when ``(W + K)`` steps would exceed ``CCMPI_REF_BUDGET_S`` (default 720 s) the arm runs as many as
fit (never fewer than 1 warm-up + 1 timed step), reports the counts it really ran in ``steps`` /
``warmup`` and says so in ``note``. Per-step time on the CPU is stable to a few per cent, so the
tokens/s of a shortened run is the same measurement with a wider error bar, not a different one.
"""
from __future__ import annotations

import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def usable_cpus() -> int:
    """CPUs this process may really use: the affinity mask, capped by a cgroup CPU quota when there is one
    (a container with 16 of a host's 128 CPUs must not start 128 BLAS threads)."""
    try:
        n = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        n = os.cpu_count() or 1
    try:
        with open("/sys/fs/cgroup/cpu.max") as f:
            quota, period = f.read().split()[:2]
        if quota != "max":
            n = min(n, max(1, -(-int(quota) // int(period))))
    except (OSError, ValueError):
        pass
    return max(1, n)


def run_reference(args, emit, grid_for, cfg):
    t_start = time.perf_counter()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_root, "model", "func_impl.py")):
        # not expected (the tree is tracked); a checkout that lost it can still rebuild it from /root/reference
        import subprocess
        if rank == 0 and os.path.isdir("/root/reference"):
            subprocess.run(["bash", os.path.join(ROOT, "scripts", "install_reference.sh")], stdout=subprocess.DEVNULL,
                           stderr=subprocess.DEVNULL, check=False)
        for _ in range(120):        # other ranks wait for rank 0's install (no communicator exists yet)
            if os.path.isfile(os.path.join(ref_root, "mpi-test.py")) or not os.path.isdir("/root/reference"):
                break
            time.sleep(0.5)
    if not os.path.isfile(os.path.join(ref_root, "model", "func_impl.py")):
        if rank == 0:
            emit({"impl": "reference", "unavailable": "baseline/_ref is missing and /root/reference does not exist here"})
        return
    if world != args.gpus:
        if rank == 0:
            emit({"impl": "reference", "unavailable": f"--gpus {args.gpus} but WORLD_SIZE={world}"})
        return
    ncpu = usable_cpus()
    threads = max(1, ncpu // world)
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[k] = str(threads)
    os.environ["CCMPI_TRANSPORT"] = "host"
    os.environ["CCMPI_FORCE_SHIM"] = "1"
    # the reference's packages first; this repo only supplies the mpi4py stand-in
    sys.path.insert(0, ref_root)
    for name in ("model", "model.func_impl", "mpi_wrapper", "mpi_wrapper.comm", "data", "data.data_parallel_preprocess"):
        sys.modules.pop(name, None)
    import numpy as np
    from mpi4py import MPI
    from model.func_impl import (get_info, naive_collect_backward_output, naive_collect_backward_x,
                                 naive_collect_forward_input, naive_collect_forward_output)
    from data.data_parallel_preprocess import split_data
    from mpi_wrapper import Communicator
    import model.func_impl as _fi
    assert os.path.realpath(_fi.__file__).startswith(os.path.realpath(ref_root)), "reference modules must come from baseline/_ref"

    comm = MPI.COMM_WORLD
    mp, dp = grid_for(world)
    H, F = cfg["hidden"], cfg["ffn_mult"] * cfg["hidden"]
    tokens = cfg["seq"] * cfg["seqs"]
    b, s = cfg["seqs"], cfg["seq"]
    rng = np.random.default_rng(rank)
    layers = []
    # the reference's "naive" layer (func_impl.py docstrings): every FC layer is partitioned along
    # the output dimension, its input is gathered, its output is gathered again
    for name, din, dout in (("fc_q", H, H), ("fc_k", H, H), ("fc_v", H, H), ("fc_o", H, H), ("fc1", H, F), ("fc2", F, H)):
        fc = name if name.startswith("fc_") else "fc_q"   # get_info knows fc_q/k/v/o only; fc1/fc2 partition like fc_q
        mp_idx, dp_idx, mp_comm, dp_comm, pin, pout = get_info(comm=comm, rank=rank, mp_size=mp, dp_size=dp,
                                                               fc_layer="fc_q", in_dim=din, out_dim=dout)
        layers.append({"name": name, "w": (rng.standard_normal((din, pout), dtype=np.float32) / np.sqrt(din)),
                       "in": din, "out": dout})
    dpw = Communicator(dp_comm)
    X = rng.standard_normal((b * dp, s, H // mp), dtype=np.float32) * 0.5
    T = rng.standard_normal((b * dp, s, H // mp), dtype=np.float32) * 0.5
    x_loc, t_loc = split_data(x_train=X, y_train=T, mp_size=mp, dp_size=dp, rank=rank)
    lr = 1e-4

    def shard(y):  # the next layer's sharded input: this rank's slice of a gathered activation
        return naive_collect_backward_output(output_grad=y, mp_group_idx=mp_idx, mp_size=mp)

    def step():
        acts = []
        h = x_loc
        # forward: q, k, v from the same input; o; fc1; fc2 -- each gather-in / GEMM / gather-out
        def fc_fwd(layer, x_sh):
            x_full = naive_collect_forward_input(x=x_sh, mp_comm=mp_comm, mp_size=mp)
            y_loc = x_full @ layer["w"]
            y_full = naive_collect_forward_output(out=y_loc, mp_comm=mp_comm, mp_size=mp)
            acts.append((layer, x_full))
            return y_full
        q, k, v = (fc_fwd(layers[i], h) for i in range(3))
        a = v * (1.0 / (1.0 + np.exp(-(q * k))))
        o = fc_fwd(layers[3], shard(a))
        h1 = naive_collect_forward_input(x=h, mp_comm=mp_comm, mp_size=mp) + o
        u = fc_fwd(layers[4], shard(h1))
        g = np.tanh(0.7978845608 * (u + 0.044715 * u ** 3))
        act = 0.5 * u * (1.0 + g)
        y = h1 + fc_fwd(layers[5], shard(act))
        t_full = naive_collect_forward_input(x=t_loc, mp_comm=mp_comm, mp_size=mp)
        diff = y - t_full
        loss = float(np.mean(diff * diff))
        # backward (weight + input gradients of every FC layer)
        grads = {}

        def fc_bwd(idx, dy_full):
            layer, x_full = acts[idx]
            dy_loc = naive_collect_backward_output(output_grad=dy_full, mp_group_idx=mp_idx, mp_size=mp)
            grads[idx] = np.tensordot(x_full, dy_loc, axes=([0, 1], [0, 1]))
            dx_part = dy_loc @ layer["w"].T
            dx_sh = naive_collect_backward_x(grad_x=np.ascontiguousarray(dx_part), mp_comm=mp_comm, mp_size=mp)
            return naive_collect_forward_input(x=dx_sh, mp_comm=mp_comm, mp_size=mp)
        dy = (2.0 / diff.size) * diff
        dact = fc_bwd(5, dy)
        du = dact * (0.5 * (1.0 + g) + 0.5 * u * (1.0 - g * g) * 0.7978845608 * (1.0 + 3 * 0.044715 * u * u))
        dh1 = dy + fc_bwd(4, du)
        da = fc_bwd(3, dh1)
        sg = 1.0 / (1.0 + np.exp(-(q * k)))
        dv, dqk = da * sg, da * v * sg * (1.0 - sg)
        dh = dh1 + fc_bwd(2, dv) + fc_bwd(1, dqk * q) + fc_bwd(0, dqk * k)
        # data-parallel gradient all-reduce through the reference's Communicator, then SGD
        for idx, layer in enumerate(layers):
            gsum = np.empty_like(grads[idx])
            dpw.Allreduce(np.ascontiguousarray(grads[idx]), gsum, op=MPI.SUM)
            layer["w"] -= lr * (gsum / dp)
        return loss

    comm.Barrier()
    t0 = time.perf_counter()
    step()
    comm.Barrier()
    first = time.perf_counter() - t0
    first = max(comm.allgather(first))
    budget = float(os.environ.get("CCMPI_REF_BUDGET_S", "720"))
    warm, steps, note = max(3, args.warmup), args.steps, None
    fit = int((budget - (time.perf_counter() - t_start)) / (first * 1.05))     # steps that still fit, first one excluded
    if (warm - 1) + steps > fit:
        warm = max(1, min(warm, 1 + (fit - 1) // 4))        # keep about a fifth of the run as warm-up
        steps = max(1, min(steps, fit - (warm - 1)))
        note = (f"one CPU step takes {first:.1f} s on {threads} threads/rank: the requested {max(3, args.warmup)}+{args.steps} "
                f"steps do not fit the {budget:.0f} s wall budget (CCMPI_REF_BUDGET_S); ran {warm}+{steps}")
    for _ in range(warm - 1):
        step()
    comm.Barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    comm.Barrier()
    el = max(comm.allgather(time.perf_counter() - t0))
    if rank == 0:
        val = tokens * dp * steps / el
        ms = round(el / steps * 1e3, 2)
        emit({"impl": "reference", "metric": cfg["metric"], "value": round(val, 1), "unit": "tokens/s", "n_gpus": world,
              "steps": steps, "warmup": warm, "ms_per_step": ms, "higher_is_better": True, **({"note": note} if note else {}),
              "scaling": "weak", "vs_baseline": None, "dtype": "fp32 (NumPy on CPU; the reference has no GPU or bf16 path)",
              "data": "synthetic",
              "config": {"model": f"transformer FC layers fc_q/k/v/o + fc1/fc2, hidden {H}, ffn {cfg['ffn_mult']}x (random init)",
                         "global_batch": cfg["seqs"] * dp, "seq_len": cfg["seq"], "parallelism": f"mp{mp}xdp{dp}",
                         "threads_per_rank": threads},
              "mpi_runtime": "in-repo shared-memory stand-in for mpi4py (no MPI on the image); reference code unmodified",
              # the reference never leaves the host: its end-to-end step IS its step (inputs are NumPy arrays in
              # host memory, the loss is a Python float), so no copies are declared
              "e2e": {"value": round(val, 1), "unit": "tokens/s", "ms_per_step": ms, "h2d_bytes_per_step": 0,
                      "d2h_bytes_per_step": 0, "note": "CPU-resident: no host<->device copies exist on this arm"},
              "native_libs_mapped": sorted({os.path.basename(l.split()[-1]) for l in open("/proc/self/maps")
                                            if "libccmpi" in l}),
              "gpu_launches": 0})
