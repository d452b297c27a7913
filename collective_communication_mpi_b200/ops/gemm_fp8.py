"""fp8 GEMMs on the CTA-pair tcgen05 kernel.

* per-tensor-scaled e4m3 (``gemm_fp8``, csrc/gemm_fp8.cu, ``UTCQMMA.2CTA``);
* block-scaled MXFP8 (``kind::mxf8f6f4.block_scale``: e4m3 payload, one e8m0 scale per 32 values;
  csrc/fused_gemm_mx.cu): the plain GEMM, the CUDA quantiser, and the two fused
  GEMM+collective ops of the tensor-parallel forward (``AllGatherGemmMX``, ``GemmReduceScatterMX``).

Run on B200 in round 2 (tests/test_gpu_fp8.py, tests/test_gpu_mx_fused.py; numbers in
profiles/R2_RESULTS.md). The block-scaled configuration is BASELINE.md's config 3.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N
from .fused import FusedOp

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


# -------------------------------------------------------------------------------------------------
# MX (block-scaled) quantisation: host side of the planned tcgen05 ``kind::mxf8f6f4.block_scale`` GEMM
# (DESIGN.md section 6). Pure PyTorch, runs on CPU or GPU; the kernel that consumes it is not built yet.
# -------------------------------------------------------------------------------------------------
MX_BLOCK = 32          # elements sharing one scale (OCP MXFP8)
SF_ROWS, SF_COLS = 128, 4   # one 512-byte chunk of the hardware scale layout: 128 rows x 4 scales


def quantize_mxfp8(x: torch.Tensor):
    """``x[R, K]`` -> ``(q, sf)``: ``q`` is ``float8_e4m3fn`` ``[R, K]``, ``sf`` is ``uint8 [R, K/32]`` holding one
    UE8M0 scale (biased power-of-two exponent) per 1 x 32 block, chosen so the block's largest magnitude fits
    e4m3's range: ``x ~= q * 2**(sf - 127)``."""
    if x.dim() != 2 or x.shape[1] % MX_BLOCK:
        raise ValueError(f"expected [R, K] with K a multiple of {MX_BLOCK}, got {tuple(x.shape)}")
    R, K = x.shape
    blocks = x.float().reshape(R, K // MX_BLOCK, MX_BLOCK)
    amax = blocks.abs().amax(dim=-1).clamp_min(2.0 ** -126)
    exp = torch.ceil(torch.log2(amax / E4M3_MAX)).clamp(-127, 127)
    q = (blocks / torch.exp2(exp).unsqueeze(-1)).clamp(-E4M3_MAX, E4M3_MAX).reshape(R, K).to(torch.float8_e4m3fn)
    return q, (exp + 127).to(torch.uint8)


def dequantize_mxfp8(q: torch.Tensor, sf: torch.Tensor) -> torch.Tensor:
    R, K = q.shape
    scale = torch.exp2(sf.float() - 127.0).unsqueeze(-1)
    return (q.float().reshape(R, K // MX_BLOCK, MX_BLOCK) * scale).reshape(R, K)


def sf_chunk_offset(r: int, k: int, k_scales: int) -> int:
    """Byte offset of the scale of row ``r``, scale column ``k`` in the layout the tensor cores read
    (cutlass/detail/sm100_blockscaled_layout.hpp): 512-byte chunks of 128 rows x 4 scales, K-fastest, and
    inside a chunk ``(r % 32) * 16 + (r // 32) * 4 + k``."""
    chunks_k = (k_scales + SF_COLS - 1) // SF_COLS
    chunk = (r // SF_ROWS) * chunks_k + k // SF_COLS
    rr, kk = r % SF_ROWS, k % SF_COLS
    return chunk * SF_ROWS * SF_COLS + (rr % 32) * 16 + (rr // 32) * 4 + kk


def swizzle_scales(sf: torch.Tensor) -> torch.Tensor:
    """Row-major ``[R, K/32]`` scales -> the chunked hardware layout (flat ``uint8``, rows and scale columns
    padded to whole chunks with the neutral exponent 127)."""
    R, Ks = sf.shape
    Rp, Kp = -(-R // SF_ROWS) * SF_ROWS, -(-Ks // SF_COLS) * SF_COLS
    pad = torch.full((Rp, Kp), 127, dtype=torch.uint8, device=sf.device)
    pad[:R, :Ks] = sf
    # [row chunk, 4 x 32 rows, k chunk, 4] -> [row chunk, k chunk, 32, 4 (row quarter), 4 (k)]
    t = pad.reshape(Rp // SF_ROWS, 4, 32, Kp // SF_COLS, SF_COLS).permute(0, 3, 2, 1, 4)
    return t.contiguous().reshape(-1)


def gemm_mxfp8(a_q: torch.Tensor, a_sf: torch.Tensor, b_q: torch.Tensor, b_sf: torch.Tensor, *,
               out_dtype=torch.bfloat16, out=None, num_sms: int = 0, stream=None) -> torch.Tensor:
    """``out[M,N] = dequant(a) @ dequant(b).T`` on the block-scaled tensor-core path (csrc/gemm_mxfp8.cu).
    ``a_q [M,K]`` / ``b_q [N,K]``: ``float8_e4m3fn``; ``a_sf [M,K/32]`` / ``b_sf [N,K/32]``: row-major UE8M0
    scales as returned by ``quantize_mxfp8`` (re-laid out here), or 1-D tensors already in the hardware
    layout (``swizzle_scales``; do that once for weights)."""
    if a_q.dtype != torch.float8_e4m3fn or b_q.dtype != torch.float8_e4m3fn:
        raise TypeError("gemm_mxfp8 takes float8_e4m3fn operands")
    M, K = a_q.shape
    Nn = b_q.shape[0]
    if b_q.shape[1] != K or K % 128:
        raise ValueError("shapes: a [M,K], b [N,K], K a multiple of 128")

    def hw_layout(sf, rows):
        if sf.dim() == 1:
            if sf.numel() != -(-rows // SF_ROWS) * (K // 128) * SF_ROWS * SF_COLS:
                raise ValueError("pre-swizzled scales have the wrong size")
            return sf
        if sf.shape != (rows, K // MX_BLOCK):
            raise ValueError("row-major scales must be [rows, K/32]")
        return swizzle_scales(sf)

    sa, sb = hw_layout(a_sf, M), hw_layout(b_sf, Nn)
    if out is None:
        out = torch.empty((M, Nn), dtype=out_dtype, device=a_q.device)
    s = stream if stream is not None else torch.cuda.current_stream()
    rc = N.device_lib().ccmpi_gemm_mxfp8(a_q.data_ptr(), sa.data_ptr(), b_q.data_ptr(), sb.data_ptr(), out.data_ptr(), M, Nn, K,
                                         a_q.stride(0), b_q.stride(0), out.stride(0),
                                         N.BF16 if out.dtype == torch.bfloat16 else N.F32, num_sms, s.cuda_stream)
    N.check(rc, "ccmpi_gemm_mxfp8")
    N.count_launch()
    return out


# -------------------------------------------------------------------------------------------------
# Block-scaled GEMMs on the generic CTA-pair kernel, plain and fused with the collectives
# (csrc/fused_gemm_mx.cu): BASELINE config 3's fp8 forward products.
# -------------------------------------------------------------------------------------------------
def sf_bytes(rows: int, K: int) -> int:
    """Size of the chunked scale array of a ``[rows, K]`` operand."""
    return -(-rows // SF_ROWS) * -(-K // 128) * SF_ROWS * SF_COLS


def quantize_mxfp8_cuda(x: torch.Tensor, q: torch.Tensor = None, sf: torch.Tensor = None, stream=None):
    """bf16 ``[R, K]`` -> ``(q, sf)``: e4m3 values and UE8M0 scales ALREADY in the chunked hardware layout
    (flat uint8, ``sf_bytes(R, K)`` long), in one kernel. ``q`` / ``sf`` may be preallocated (e.g. symmetric)."""
    if x.dtype != torch.bfloat16 or x.dim() != 2 or x.stride(1) != 1 or x.shape[1] % MX_BLOCK:
        raise ValueError("quantize_mxfp8_cuda takes a bfloat16 [R, K] tensor with unit column stride, K % 32 == 0")
    R, K = x.shape
    if q is None:
        q = torch.empty((R, K), dtype=torch.uint8, device=x.device)
    if sf is None:
        sf = torch.full((sf_bytes(R, K),), 127, dtype=torch.uint8, device=x.device)
    s = stream if stream is not None else torch.cuda.current_stream()
    N.check(N.device_lib().ccmpi_quantize_mxfp8(x.data_ptr(), x.stride(0), q.data_ptr(), q.stride(0), sf.data_ptr(), R, K,
                                                s.cuda_stream), "ccmpi_quantize_mxfp8")
    N.count_launch()
    return q, sf


def _u8(t: torch.Tensor) -> torch.Tensor:
    return t.view(torch.uint8) if t.dtype != torch.uint8 else t


def gemm_mx(a_q: torch.Tensor, a_sf: torch.Tensor, b_q: torch.Tensor, b_sf: torch.Tensor, *, out_dtype=torch.bfloat16,
            out=None, num_sms: int = 0, stream=None) -> torch.Tensor:
    """``dequant(a) @ dequant(b).T`` on the generic CTA-pair kernel's block-scaled mainloop. ``a_q [M,K]`` /
    ``b_q [N,K]``: e4m3 (or their uint8 views); ``a_sf`` / ``b_sf``: chunked scale arrays (flat uint8)."""
    a_q, b_q = _u8(a_q), _u8(b_q)
    M, K = a_q.shape
    Nn = b_q.shape[0]
    if out is None:
        out = torch.empty((M, Nn), dtype=out_dtype, device=a_q.device)
    s = stream if stream is not None else torch.cuda.current_stream()
    N.check(N.device_lib().ccmpi_gemm_mx(a_q.data_ptr(), a_sf.data_ptr(), b_q.data_ptr(), b_sf.data_ptr(), out.data_ptr(), M, Nn, K,
                                         a_q.stride(0), b_q.stride(0), out.stride(0),
                                         N.BF16 if out.dtype == torch.bfloat16 else N.F32, num_sms, s.cuda_stream), "ccmpi_gemm_mx")
    N.count_launch()
    return out


def _pp(tensors):
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def ag_gemm_mx_raw(comms, xq_off: int, ldx: int, sfl_off: int, xfull_off: int, sffull_off: int, B, sfb, Cout, M: int, Nn: int,
                   K: int, *, flag_base: int, epoch: int, flag_target: int = 0, trace=None, num_sms: int = 0, stream=None):
    from .collectives import _arr, _sp
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_ag_gemm_mx(a, n, xq_off, ldx, sfl_off, xfull_off, sffull_off, _pp(B), _pp(sfb), _pp(Cout), M, Nn,
                                               K, B[0].stride(0), Cout[0].stride(0),
                                               N.BF16 if Cout[0].dtype == torch.bfloat16 else N.F32, flag_base, epoch,
                                               flag_target & 0xFFFFFFFF, C.c_void_p(trace.data_ptr() if trace is not None else None),
                                               num_sms, _sp(stream))
    N.check(rc, "ccmpi_fused_ag_gemm_mx")
    N.count_launch()


def gemm_rs_mx_raw(comms, A, sfa, B, sfb, slots_off: int, Cout, M: int, Nn: int, K: int, *, flag_base: int, epoch: int,
                   residual=None, trace=None, num_sms: int = 0, stream=None):
    from .collectives import _arr, _sp
    a, n = _arr(comms)
    rc = N.device_lib().ccmpi_fused_gemm_rs_mx(a, n, _pp(A), _pp(sfa), _pp(B), _pp(sfb), slots_off, _pp(Cout), M, Nn, K,
                                               A[0].stride(0), B[0].stride(0), Cout[0].stride(0),
                                               N.BF16 if Cout[0].dtype == torch.bfloat16 else N.F32, flag_base, epoch,
                                               _pp(residual) if residual is not None else None,
                                               residual[0].stride(0) if residual is not None else 0,
                                               C.c_void_p(trace.data_ptr() if trace is not None else None), num_sms, _sp(stream))
    N.check(rc, "ccmpi_fused_gemm_rs_mx")
    N.count_launch()


class QuantizedWeight:
    """A weight shard quantised once for the block-scaled kernels: ``[N, K]`` e4m3 (K-major) + chunked scales.
    ``w`` is given in the reference's ``[in, out]`` layout (``b_kn``) or as ``[out, in]``."""

    def __init__(self, w: torch.Tensor, *, b_kn: bool = True):
        wt = (w.t() if b_kn else w).contiguous().to(torch.bfloat16)
        self.q, self.sf = quantize_mxfp8_cuda(wt)
        self.N, self.K = wt.shape


class AllGatherGemmMX(FusedOp):
    """``gather_last_axis(x) @ w`` with block-scaled fp8 operands: every rank quantises its ``[M, K/n]`` shard,
    and the fused kernel pushes e4m3 payload + scale chunks -- half the NVLink bytes of the bf16 gather -- while
    its tensor cores start on the local K range (func_impl.py:76-91 + the FC GEMM, in MXFP8)."""

    def __init__(self, comm, M: int, K: int):
        from .fused import _ag_flag_words, _ctx
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        if K % n or (K // n) % 128:
            raise ValueError(f"K/n = {K}/{n} must be a multiple of 128")
        self.M, self.K, self.chunk = M, K, K // n
        self.xq_off, self.xq = h.empty((M, self.chunk), torch.uint8)
        self.sfl_off, self.sfl = h.empty((sf_bytes(M, self.chunk),), torch.uint8)
        self.sfl.fill_(127)
        self.xfull = [h.empty((M, K), torch.uint8) for _ in range(2)]
        self.sffull = [h.empty((sf_bytes(M, K),), torch.uint8) for _ in range(2)]
        for _, t in self.sffull:
            t.fill_(127)
        self.flag_base = self.ctx.new_slot(_ag_flag_words(M, n))
        self.epoch, self.flag_total = 0, 0

    def __call__(self, x: torch.Tensor, w: QuantizedWeight, *, out=None, out_dtype=torch.bfloat16, num_sms: int = 0, stream=None):
        self._check_open()
        from ..utils import ktrace as KT
        from .fused import ag_ctas
        quantize_mxfp8_cuda(x.reshape(self.M, self.chunk), self.xq, self.sfl, stream=stream)
        if out is None:
            out = torch.empty((self.M, w.N), dtype=out_dtype, device=x.device)
        self.epoch += 1
        self.flag_total += ag_ctas(2, self.M, w.N, 1, num_sms)
        par = self.epoch & 1
        ag_gemm_mx_raw(self.ctx.dev.comms, self.xq_off, self.chunk, self.sfl_off, self.xfull[par][0], self.sffull[par][0],
                       [w.q], [w.sf], [out], self.M, w.N, self.K, flag_base=self.flag_base, epoch=self.epoch,
                       flag_target=self.flag_total, trace=KT.buffer_for(self, "ag_gemm_mx"), num_sms=num_sms, stream=stream)
        return out


class GemmReduceScatterMX(FusedOp):
    """``reduce_scatter_last_axis(a @ w)`` with block-scaled fp8 operands and bf16 partial tiles
    (the GEMM + func_impl.py:150-187 in MXFP8). ``a``: bf16 ``[M, K]`` (quantised here) ."""

    def __init__(self, comm, M: int, Nfull: int, *, dtype=torch.bfloat16):
        from .fused import _ctx
        self.ctx = _ctx(comm)
        h, n = self.ctx.heap, self.ctx.n
        if Nfull % n or (Nfull // n) % 128:
            raise ValueError(f"N/n = {Nfull}/{n} must be a multiple of 128")
        self.M, self.N, self.chunk, self.dtype = M, Nfull, Nfull // n, dtype
        self.slots = [h.empty((n, M, self.chunk), dtype) for _ in range(2)]
        self.flag_base = self.ctx.new_slot(((M + 127) // 128) * (self.chunk // 128) * n)
        self.epoch = 0

    def __call__(self, a: torch.Tensor, w: QuantizedWeight, *, out=None, add=None, num_sms: int = 0, stream=None):
        self._check_open()
        from ..utils import ktrace as KT
        aq, asf = quantize_mxfp8_cuda(a.reshape(self.M, -1), stream=stream)
        if out is None:
            out = torch.empty((self.M, self.chunk), dtype=self.dtype, device=a.device)
        self.epoch += 1
        gemm_rs_mx_raw(self.ctx.dev.comms, [aq], [asf], [w.q], [w.sf], self.slots[self.epoch & 1][0], [out], self.M, self.N,
                       aq.shape[1], flag_base=self.flag_base, epoch=self.epoch, residual=[add] if add is not None else None,
                       trace=KT.buffer_for(self, "gemm_rs_mx"), num_sms=num_sms, stream=stream)
        return out
