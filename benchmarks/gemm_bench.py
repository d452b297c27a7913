"""tcgen05 GEMM vs cuBLAS (torch.matmul) on the FC shapes of the benchmark configurations.
One JSON line per shape to stdout and gpurun_out/gemm_bench.jsonl. L2 is flushed between timed
iterations by writing a 256 MiB buffer."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from collective_communication_mpi_b200.ops.gemm import gemm  # noqa: E402

SHAPES = [  # (name, M, N, K, a_km, b_kn)
    ("square_8k", 8192, 8192, 8192, False, False),
    ("fwd_qkv_mp2  y=xW", 8192, 6144, 4096, False, True),
    ("fwd_fc1_mp2  y=xW", 8192, 8192, 4096, False, True),
    ("fwd_fc2_mp2  y=xW", 8192, 4096, 8192, False, True),
    ("fwd_o_mp2    y=xW", 8192, 4096, 2048, False, True),
    ("dx_fc1_mp2   dyW^T", 8192, 4096, 8192, False, False),
    ("dw_fc1_mp2   x^Tdy", 4096, 8192, 8192, True, True),
    ("dw_o_mp2     x^Tdy", 2048, 4096, 8192, True, True),
    ("fwd_qkv_mp1", 8192, 12288, 4096, False, True),
    ("fwd_fc1_mp1", 8192, 16384, 4096, False, True),
    ("dw_fc1_mp1", 4096, 16384, 8192, True, True),
]


def bench(fn, flush, iters=12):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[0], ts[len(ts) // 2]


def main():
    torch.manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for name, M, N, K, a_km, b_kn in SHAPES:
        A = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
        B = (torch.randn(K, N, device="cuda") * 0.5).bfloat16()
        a = A.t().contiguous() if a_km else A
        b = B if b_kn else B.t().contiguous()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        ours = lambda: gemm(a, b, a_km=a_km, b_kn=b_kn, out=out)  # noqa: E731
        # cuBLAS gets the same storage: transposes are views, never materialised
        ta = a.t() if a_km else a
        tb = b if b_kn else b.t()
        ref_out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        cublas = lambda: torch.matmul(ta, tb, out=ref_out)  # noqa: E731
        ours()
        cublas()
        torch.cuda.synchronize()
        err = (out.float() - ref_out.float()).abs().max().item()
        o_best, o_med = bench(ours, flush)
        c_best, c_med = bench(cublas, flush)
        fl = 2.0 * M * N * K
        row = {"shape": name, "M": M, "N": N, "K": K, "ours_tflops_best": round(fl / o_best / 1e9, 1),
               "ours_tflops_median": round(fl / o_med / 1e9, 1), "cublas_tflops_best": round(fl / c_best / 1e9, 1),
               "cublas_tflops_median": round(fl / c_med / 1e9, 1), "ratio_median": round(c_med / o_med, 3),
               "max_abs_diff_vs_cublas": err, "mode": "1cta" if os.environ.get("CCMPI_GEMM_1CTA") == "1" else "2cta"}
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    with open(os.path.join("gpurun_out", f"gemm_bench_{rows[0]['mode']}.jsonl"), "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
