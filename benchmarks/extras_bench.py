"""Single-GPU micro-benchmarks of the kernels that were written after round 1's GPU budget was spent:
fused elementwise ops vs the eager PyTorch chains they replace, SGD-in-the-dW-epilogue vs GEMM + optimizer
pass, and the fp8 / block-scaled fp8 GEMMs vs the bf16 GEMM. Every case is checked against its fp32
reference before it is timed. CUDA events, median of 10, L2 flushed between iterations.

    python benchmarks/extras_bench.py [--tokens 8192 --hidden 4096]

One JSON line per case. Nothing here has run on hardware yet; a failing check aborts that case only.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

BF = torch.bfloat16


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=8192)
    ap.add_argument("--hidden", type=int, default=4096)
    args = ap.parse_args()
    T, H = args.tokens, args.hidden
    from collective_communication_mpi_b200.ops import elementwise as EW
    from collective_communication_mpi_b200.ops import gemm_fp8 as Q
    from collective_communication_mpi_b200.ops.gemm import gemm, gemm_sgd_, matmul_tn

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timeit(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
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
        return ts[len(ts) // 2]

    def rnd(*shape, seed, scale=0.5):
        g = torch.Generator(device="cuda").manual_seed(seed)
        return (torch.randn(*shape, device="cuda", generator=g) * scale).to(BF)

    def close(a, b, tol):
        return (a.float() - b.float()).abs().max().item() <= tol * max(1.0, b.float().abs().max().item())

    def case(name, check, ours, baseline, bytes_moved=None, flops=None):
        try:
            ok = bool(check())
        except Exception as e:  # noqa: BLE001
            print(json.dumps({"case": name, "error": repr(e)[:300]}), flush=True)
            return
        if not ok:
            print(json.dumps({"case": name, "error": "numerics check failed"}), flush=True)
            return
        t_o, t_b = timeit(ours), timeit(baseline)
        row = {"case": name, "ours_ms": round(t_o, 4), "baseline_ms": round(t_b, 4), "speedup": round(t_b / t_o, 3)}
        if bytes_moved:
            row["ours_GBs"] = round(bytes_moved / t_o / 1e6, 1)
        if flops:
            row["ours_tflops"] = round(flops / t_o / 1e9, 1)
            row["baseline_tflops"] = round(flops / t_b / 1e9, 1)
        print(json.dumps(row), flush=True)

    # ---- elementwise -----------------------------------------------------------------------------
    qkv, da = rnd(T, 3 * H, seed=1), rnd(T, H, seed=2)

    def eager_gate(x):
        q, k, v = x.chunk(3, dim=-1)
        return v * torch.sigmoid(q * k)

    def eager_gate_bwd():
        x = qkv.detach().requires_grad_(True)
        eager_gate(x).backward(da)
        return x.grad

    case("gate fwd", lambda: close(EW.gate_forward(qkv), EW.gate_reference(qkv.float()), 2e-2),
         lambda: EW.gate_forward(qkv), lambda: eager_gate(qkv), bytes_moved=T * H * 2 * 4)
    case("gate fwd+bwd", lambda: close(EW.gate_backward(da, qkv), EW.gate_backward_reference(da.float(), qkv.float()), 2e-2),
         lambda: (EW.gate_forward(qkv), EW.gate_backward(da, qkv)), eager_gate_bwd, bytes_moved=T * H * 2 * 11)
    u, du = rnd(T, 4 * H, seed=3, scale=2.0), rnd(T, 4 * H, seed=4)

    def eager_gelu_bwd():
        x = u.detach().requires_grad_(True)
        torch.nn.functional.gelu(x, approximate="tanh").backward(du)
        return x.grad

    case("gelu fwd", lambda: close(EW.gelu_forward(u), EW.gelu_reference(u.float()), 2e-2),
         lambda: EW.gelu_forward(u), lambda: torch.nn.functional.gelu(u, approximate="tanh"), bytes_moved=T * 4 * H * 2 * 2)
    case("gelu fwd+bwd", lambda: close(EW.gelu_backward(du, u), EW.gelu_backward_reference(du.float(), u.float()), 2e-2),
         lambda: (EW.gelu_forward(u), EW.gelu_backward(du, u)), eager_gelu_bwd, bytes_moved=T * 4 * H * 2 * 5)
    y, t = rnd(T, H, seed=5), rnd(T, H, seed=6)

    def eager_mse():
        yy = y.detach().requires_grad_(True)
        loss = torch.nn.functional.mse_loss(yy.float(), t.float())
        loss.backward()
        return loss, yy.grad

    def mse_ok():
        loss, dy = EW.mse_loss_and_grad(y, t)
        rl, rdy = EW.mse_reference(y.float(), t.float())
        return abs(float(loss) - float(rl)) < 1e-4 * max(1.0, float(rl)) and close(dy, rdy, 2e-2)

    case("mse loss+grad", mse_ok, lambda: EW.mse_loss_and_grad(y, t), eager_mse, bytes_moved=T * H * 2 * 3)

    # ---- SGD in the dW epilogue --------------------------------------------------------------------
    x, dy = rnd(T, H, seed=7), rnd(T, 4 * H, seed=8)
    w = rnd(H, 4 * H, seed=9, scale=0.02)
    lr = 1e-3

    def sgd_ok():
        w0 = w.clone()
        ref = w0.float() - lr * (x.float().t() @ dy.float())
        gemm_sgd_(w0, x, dy, lr, a_km=True, b_kn=True)
        return close(w0, ref, 1e-2)

    grad = torch.empty(H, 4 * H, device="cuda", dtype=BF)

    def unfused():
        matmul_tn(x, dy, out=grad)
        torch._foreach_add_([w], [grad], alpha=-lr)

    case("dW + SGD (fc1 shape)", sgd_ok, lambda: gemm_sgd_(w, x, dy, lr, a_km=True, b_kn=True), unfused, flops=2.0 * T * H * 4 * H)

    # ---- fp8 GEMMs (vs our bf16 GEMM of the same shape) --------------------------------------------
    a, b = rnd(T, H, seed=10, scale=1.0), rnd(4 * H, H, seed=11, scale=1.0)
    out_bf = torch.empty(T, 4 * H, device="cuda", dtype=BF)
    aq, sa = Q.quantize_e4m3(a)
    bq, sb = Q.quantize_e4m3(b)
    case("fp8 per-tensor GEMM (fc1 fwd)",
         lambda: close(Q.gemm_fp8(aq, bq, float(sa * sb)), (aq.float() @ bq.float().t()) * (sa * sb), 2e-2),
         lambda: Q.gemm_fp8(aq, bq, float(sa * sb), out=out_bf), lambda: gemm(a, b, out=out_bf), flops=2.0 * T * H * 4 * H)
    am, asf = Q.quantize_mxfp8(a)
    bm, bsf = Q.quantize_mxfp8(b)
    asw, bsw = Q.swizzle_scales(asf), Q.swizzle_scales(bsf)          # hardware layout, produced once
    case("mxfp8 block-scaled GEMM (fc1 fwd)",
         lambda: close(Q.gemm_mxfp8(am, asf, bm, bsf), Q.dequantize_mxfp8(am, asf) @ Q.dequantize_mxfp8(bm, bsf).t(), 2e-2),
         lambda: Q.gemm_mxfp8(am, asw, bm, bsw, out=out_bf), lambda: gemm(a, b, out=out_bf), flops=2.0 * T * H * 4 * H)


if __name__ == "__main__":
    main()
