"""fp8 (e4m3) GEMM on the CTA-pair tcgen05 kernel (csrc/gemm_fp8.cu) with per-tensor scales.

EXPERIMENTAL (round 1): compiled and SASS-checked (``UTCQMMA.2CTA``) but not yet run on a GPU --
the round's GPU budget was spent when it was written. Not used by any default path. First step
towards the block-scaled fp8 configuration of BASELINE.md (config 3).
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N

E4M3_MAX = 448.0


def quantize_e4m3(x: torch.Tensor):
    """Per-tensor symmetric quantisation: returns ``(x_fp8, scale)`` with ``x ~= x_fp8 * scale``."""
    amax = x.detach().abs().amax().float().clamp_min(1e-12)
    scale = amax / E4M3_MAX
    return (x.float() / scale).to(torch.float8_e4m3fn), scale


def gemm_fp8(a: torch.Tensor, b: torch.Tensor, scale: float = 1.0, *, out_dtype=torch.bfloat16, out=None,
             num_sms: int = 0, stream=None) -> torch.Tensor:
    """``out[M,N] = scale * a[M,K] @ b[N,K].T`` for ``float8_e4m3fn`` operands (both K-major)."""
    if a.dtype != torch.float8_e4m3fn or b.dtype != torch.float8_e4m3fn:
        raise TypeError("gemm_fp8 takes float8_e4m3fn operands")
    if a.dim() != 2 or b.dim() != 2 or a.shape[1] != b.shape[1] or a.stride(1) != 1 or b.stride(1) != 1:
        raise ValueError("operands must be [M,K] and [N,K] with unit stride on K")
    M, K = a.shape
    Nn = b.shape[0]
    if out is None:
        out = torch.empty((M, Nn), dtype=out_dtype, device=a.device)
    lib = N.device_lib()
    fn = lib.ccmpi_gemm_fp8_e4m3
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p] * 3 + [C.c_int] * 6 + [C.c_float, C.c_int, C.c_int, C.c_void_p]
    s = stream if stream is not None else torch.cuda.current_stream()
    rc = fn(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, Nn, K, a.stride(0), b.stride(0), out.stride(0),
            float(scale), N.BF16 if out.dtype == torch.bfloat16 else N.F32, num_sms, s.cuda_stream)
    N.check(rc, "ccmpi_gemm_fp8_e4m3")
    N.count_launch()
    return out


def linear_fp8(x: torch.Tensor, w_nk: torch.Tensor, **kw) -> torch.Tensor:
    """bf16 in / bf16 out convenience: quantise both operands per tensor, multiply in fp8."""
    xq, sx = quantize_e4m3(x)
    wq, sw = quantize_e4m3(w_nk)
    return gemm_fp8(xq, wq, float(sx * sw), **kw)
