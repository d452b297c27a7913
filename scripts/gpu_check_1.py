"""First-contact GPU check (1 GPU): tcgen05 GEMM in all operand layouts + emulated-rank
collectives. Every case is isolated in a subprocess so a trap in one kernel does not hide the
others; results go to gpurun_out/check1.jsonl.

    python scripts/gpu_check_1.py            # driver: runs all cases
    python scripts/gpu_check_1.py CASE       # one case in-process
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "gpurun_out")


def case_gemm(name, M, N, K, a_km, b_kn, out_dtype, accumulate=False):
    import torch
    from collective_communication_mpi_b200.ops.gemm import gemm
    torch.manual_seed(0)
    dev = "cuda:0"
    A = (torch.randn(M, K, device=dev) * 0.5).to(torch.bfloat16)
    B = (torch.randn(K, N, device=dev) * 0.5).to(torch.bfloat16)
    a = A.t().contiguous() if a_km else A
    b = B if b_kn else B.t().contiguous()
    ref = A.float() @ B.float()
    od = torch.bfloat16 if out_dtype == "bf16" else torch.float32
    out = None
    if accumulate:
        out = torch.ones(M, N, device=dev, dtype=torch.float32)
        ref = ref + 1.0
    c = gemm(a, b, a_km=a_km, b_kn=b_kn, out=out, out_dtype=od, accumulate=accumulate)
    torch.cuda.synchronize()
    err = (c.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    rel = err / max(scale, 1e-6)
    tol = 2e-2 if out_dtype == "bf16" else 2e-3
    res = {"case": name, "max_abs_err": err, "ref_max": scale, "rel": rel, "ok": bool(rel < tol)}
    if M * N * K >= 2048 ** 3:
        # timing: inputs > L2? 8192^2 bf16 = 128 MiB each, yes for the big case
        for _ in range(3):
            gemm(a, b, a_km=a_km, b_kn=b_kn, out=c)
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            gemm(a, b, a_km=a_km, b_kn=b_kn, out=c)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = min(ts)
        res["ms_best"] = ms
        res["tflops_best"] = 2.0 * M * N * K / ms / 1e9
        ts = []
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(A, B)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        res["cublas_tflops_best"] = 2.0 * M * N * K / min(ts) / 1e9
    return res


def _emu(nr, nbytes=64 << 20):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    heap = SymmetricHeap.create_emulated(nr, nbytes)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    return heap, comms


def case_coll(name, nr, dtype_name, op_name, count, algo_name):
    import torch
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    dt = getattr(torch, dtype_name)
    heap, comms = _emu(nr)
    es = torch.empty((), dtype=dt).element_size()
    src = heap.alloc(count * es)
    dst = heap.alloc(count * es * nr)
    tmp = heap.alloc(count * es)
    torch.manual_seed(1)
    hosts = []
    for r in range(nr):
        if dt.is_floating_point:
            x = (torch.randn(count) * 4).to(dt)
        else:
            x = torch.randint(-1000, 1000, (count,), dtype=dt)
        hosts.append(x)
        heap.view(src, (count,), dt, rank=r).copy_(x)
    op = {"sum": N.SUM, "min": N.MIN, "max": N.MAX}[op_name]
    algo = {"auto": N.AR_AUTO, "oneshot": N.AR_ONESHOT, "twoshot": N.AR_TWOSHOT}[algo_name]
    stack = torch.stack([h.double() if dt.is_floating_point else h for h in hosts])
    if op_name == "sum":
        ref = stack.sum(0)
    elif op_name == "min":
        ref = stack.min(0).values
    else:
        ref = stack.max(0).values
    ok = True
    worst = 0.0
    for it in range(3):  # repeated launches exercise epoch reuse
        CO.allreduce(comms, CO.cc_dtype(dt), op, src, dst, count, tmp_off=tmp, algo=algo)
        torch.cuda.synchronize()
        for r in range(nr):
            got = heap.view(dst, (count,), dt, rank=r).cpu()
            if dt.is_floating_point:
                e = (got.double() - ref).abs().max().item()
                tol = {torch.float32: 1e-4, torch.float64: 1e-9, torch.bfloat16: 0.25, torch.float16: 0.05}[dt]
                worst = max(worst, e)
                ok = ok and e <= tol
            else:
                ok = ok and bool((got == ref.to(dt)).all())
    heap.close()
    return {"case": name, "ok": bool(ok), "max_abs_err": worst}


def case_ag_rs_a2a(name, nr):
    import torch
    from collective_communication_mpi_b200 import _native as N
    from collective_communication_mpi_b200.ops import collectives as CO
    heap, comms = _emu(nr)
    ok = True
    # --- flat all-gather, int64 x 2 per rank (mpi-test.py:105-111 shape) and a big bf16 one
    for dt, per in ((torch.int64, 2), (torch.bfloat16, 4096 * 3), (torch.float32, 5)):
        es = torch.empty((), dtype=dt).element_size()
        s = heap.alloc(per * es)
        d = heap.alloc(per * es * nr)
        parts = []
        for r in range(nr):
            x = torch.arange(per, dtype=torch.float32).add(100 * r).to(dt)
            parts.append(x)
            heap.view(s, (per,), dt, rank=r).copy_(x)
        for _ in range(2):
            CO.allgather(comms, s, d, 1, per * es, 0, 0)
        torch.cuda.synchronize()
        ref = torch.cat(parts)
        for r in range(nr):
            ok = ok and bool((heap.view(d, (per * nr,), dt, rank=r).cpu() == ref).all())
    # --- last-axis all-gather (b, s, part) -> (b, s, part*nr) as in func_impl.py:89-90, fp64
    b_, s_, part = 4, 8, 2
    dt = torch.float64
    full = torch.arange(b_ * s_ * part * nr, dtype=dt).reshape(b_, s_, part * nr)
    so = heap.alloc(b_ * s_ * part * 8)
    do = heap.alloc(b_ * s_ * part * nr * 8)
    for r in range(nr):
        heap.view(so, (b_, s_, part), dt, rank=r).copy_(full[:, :, r * part:(r + 1) * part])
    CO.allgather(comms, so, do, b_ * s_, part * 8, part * 8, part * nr * 8)
    torch.cuda.synchronize()
    for r in range(nr):
        ok = ok and bool((heap.view(do, (b_, s_, part * nr), dt, rank=r).cpu() == full).all())
    ag_ok = ok
    # --- reduce-scatter flat (16 -> 2 int64 MIN, mpi-test.py:121-127) and last-axis fp64 SUM
    per = 2
    s = heap.alloc(per * nr * 8)
    d = heap.alloc(per * 8)
    srcs = []
    for r in range(nr):
        x = torch.randint(0, 100, (per * nr,), dtype=torch.int64)
        srcs.append(x)
        heap.view(s, (per * nr,), torch.int64, rank=r).copy_(x)
    CO.reduce_scatter(comms, N.I64, N.MIN, s, d, 1, per, 0, 0)
    torch.cuda.synchronize()
    ref = torch.stack(srcs).min(0).values
    for r in range(nr):
        ok = ok and bool((heap.view(d, (per,), torch.int64, rank=r).cpu() == ref[r * per:(r + 1) * per]).all())
    rows, width = 3, 2 * nr
    s = heap.alloc(rows * width * 8)
    d = heap.alloc(rows * (width // nr) * 8)
    g = torch.arange(nr * rows * width, dtype=torch.float64).reshape(nr, rows, width)
    for r in range(nr):
        heap.view(s, (rows, width), torch.float64, rank=r).copy_(g[r])
    CO.reduce_scatter(comms, N.F64, N.SUM, s, d, rows, width // nr, width * 8, (width // nr) * 8)
    torch.cuda.synchronize()
    tot = g.sum(0)
    w = width // nr
    for r in range(nr):
        ok = ok and bool((heap.view(d, (rows, w), torch.float64, rank=r).cpu() == tot[:, r * w:(r + 1) * w]).all())
    # bf16 vector path
    rows, w = 64, 256
    s = heap.alloc(rows * w * nr * 2)
    d = heap.alloc(rows * w * 2)
    g = torch.randn(nr, rows, w * nr).to(torch.bfloat16)
    for r in range(nr):
        heap.view(s, (rows, w * nr), torch.bfloat16, rank=r).copy_(g[r])
    CO.reduce_scatter(comms, N.BF16, N.SUM, s, d, rows, w, w * nr * 2, w * 2)
    torch.cuda.synchronize()
    tot = g.float().sum(0)
    for r in range(nr):
        e = (heap.view(d, (rows, w), torch.bfloat16, rank=r).cpu().float() - tot[:, r * w:(r + 1) * w]).abs().max().item()
        ok = ok and e < 0.1
    rs_ok = ok
    # --- all-to-all: 1 int64 per peer (mpi-test.py:164-174), and 4 KiB segments, both schedules
    for seg_elems, stag in ((1, False), (1, True), (512, False), (512, True)):
        s = heap.alloc(seg_elems * nr * 8)
        d = heap.alloc(seg_elems * nr * 8)
        for r in range(nr):
            x = torch.arange(seg_elems * nr, dtype=torch.int64) + r * 100000
            heap.view(s, (seg_elems * nr,), torch.int64, rank=r).copy_(x)
        CO.alltoall(comms, s, d, seg_elems * 8, staggered=stag)
        torch.cuda.synchronize()
        for r in range(nr):
            got = heap.view(d, (nr, seg_elems), torch.int64, rank=r).cpu()
            for j in range(nr):
                exp = torch.arange(r * seg_elems, (r + 1) * seg_elems, dtype=torch.int64) + j * 100000
                ok = ok and bool((got[j] == exp).all())
    CO.barrier(comms)
    torch.cuda.synchronize()
    heap.close()
    return {"case": name, "ok": bool(ok), "ag_ok": bool(ag_ok), "rs_ok": bool(rs_ok)}


CASES = {}
for (nm, M, N_, K, akm, bkn, od) in [
    ("gemm_kk_256", 256, 256, 128, False, False, "bf16"),
    ("gemm_kk_odd", 384, 328, 200, False, False, "bf16"),
    ("gemm_kn_1k", 1024, 1024, 1024, False, True, "bf16"),
    ("gemm_mk_1k", 1024, 1024, 1024, True, False, "bf16"),
    ("gemm_mn_1k", 1024, 1024, 1024, True, True, "f32"),
    ("gemm_kk_f32", 512, 768, 512, False, False, "f32"),
    ("gemm_kk_n128", 4096, 128, 512, False, False, "bf16"),
    ("gemm_kk_big", 8192, 8192, 8192, False, False, "bf16"),
    ("gemm_kn_cfg2", 8192, 2048, 4096, False, True, "bf16"),
    ("gemm_mn_cfg2", 4096, 2048, 8192, True, True, "f32"),
]:
    CASES[nm] = (case_gemm, (nm, M, N_, K, akm, bkn, od))
CASES["gemm_acc"] = (case_gemm, ("gemm_acc", 512, 512, 256, False, False, "f32", True))
for nr in (2, 4):
    for dtn, opn, cnt, algo in [
        ("int64", "min", 100, "oneshot"), ("int64", "sum", 10, "oneshot"), ("float32", "sum", 1 << 16, "oneshot"),
        ("float32", "sum", 1 << 18, "twoshot"), ("bfloat16", "sum", 1 << 18, "twoshot"), ("float64", "max", 4096, "twoshot"),
        ("bfloat16", "sum", 1000, "auto"), ("int32", "max", 7, "auto"),
    ]:
        nm = f"ar{nr}_{dtn}_{opn}_{cnt}_{algo}"
        CASES[nm] = (case_coll, (nm, nr, dtn, opn, cnt, algo))
    CASES[f"ag_rs_a2a_{nr}"] = (case_ag_rs_a2a, (f"ag_rs_a2a_{nr}", nr))


def run_cases(names):
    for nm in names:
        fn, args = CASES[nm]
        t0 = time.time()
        try:
            res = fn(*args)
        except Exception as e:  # noqa: BLE001 - report and let the driver isolate the rest
            res = {"case": nm, "ok": False, "error": repr(e)[:1500]}
            print("RESULT " + json.dumps(res), flush=True)
            raise
        res["secs"] = round(time.time() - t0, 2)
        print("RESULT " + json.dumps(res), flush=True)


def main():
    os.makedirs(OUT, exist_ok=True)
    if len(sys.argv) > 1:
        run_cases(sys.argv[1].split(","))
        return
    path = os.path.join(OUT, "check1.jsonl")
    results = {}

    def spawn(names, timeout):
        try:
            r = subprocess.run([sys.executable, __file__, ",".join(names)], capture_output=True, text=True, timeout=timeout)
            out, err, rc = r.stdout, r.stderr, r.returncode
        except subprocess.TimeoutExpired as e:
            out = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
            err, rc = "TIMEOUT", -9
        for l in out.splitlines():
            if l.startswith("RESULT "):
                res = json.loads(l[7:])
                results[res["case"]] = res
        return rc, err

    names = list(CASES)
    groups = [[n for n in names if n.startswith("gemm")], [n for n in names if not n.startswith("gemm")]]
    for g in groups:
        todo = list(g)
        first = True
        while todo:
            batch = todo if first else todo[:1]
            rc, err = spawn(batch, 300 if first else 120)
            done = [n for n in batch if n in results]
            missing = [n for n in batch if n not in results]
            if missing and (not first or not done):
                # the first unreported case is the one that died
                results[missing[0]] = {"case": missing[0], "ok": False, "rc": rc, "stderr": err[-1500:]}
            elif missing and first:
                # a case reported failure by exception or the process died after `done`
                last = missing[0]
                results[last] = {"case": last, "ok": False, "rc": rc, "stderr": err[-1500:]}
            todo = [n for n in todo if n not in results]
            first = False
    n_ok = 0
    with open(path, "w") as f:
        for nm in names:
            res = results.get(nm, {"case": nm, "ok": False, "missing": True})
            n_ok += bool(res.get("ok"))
            f.write(json.dumps(res) + "\n")
            print(json.dumps(res)[:600], flush=True)
    print(f"SUMMARY {n_ok}/{len(names)} ok")


if __name__ == "__main__":
    main()
