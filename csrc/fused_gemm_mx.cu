// Block-scaled fp8 (OCP MXFP8: e4m3 values, one UE8M0 scale per 1 x 32 block) GEMMs on the generic CTA-pair
// kernel -- and therefore fused with the collectives, because every Policy of fused_policies.cuh applies to
// the block-scaled mainloop unchanged (gemm_kernel_2cta.cuh, template parameter MX):
//
//   ccmpi_gemm_mx            C = dq(A) dq(B)^T                         (PlainPolicy; cross-check of the mainloop)
//   ccmpi_fused_ag_gemm_mx   C = [dq(x_0) | ... | dq(x_{n-1})] dq(B)^T (AgKPolicy): every rank pushes its e4m3
//                            shard AND its scale chunks -- half the NVLink bytes of the bf16 gather -- into the
//                            peers' staging while the tensor cores work on the local K range
//   ccmpi_fused_gemm_rs_mx   reduce-scatter of dq(A) dq(B)^T on N       (RsNPolicy): fp8 operands, bf16 partials
//   ccmpi_quantize_mxfp8     bf16 [R][K] -> e4m3 [R][K] + scales in the 512-byte-chunk layout the tensor cores read
//
// This is BASELINE config 3's "block-scaled fp8 GEMMs fused with the collectives" for the forward products of
// the tensor-parallel FC layers (func_impl.py:76-91 gather + GEMM, :150-187 GEMM + reduce-scatter); the reference
// itself has neither a GEMM nor a reduced-precision path. Operands are K-major: A [M][K], B [N][K].
#include <cuda_fp8.h>

#include "fused_policies.cuh"
#include "gemm_host.h"

using namespace gemm;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// scale buffer as a [chunks][128] matrix of 32-bit words, one 512-byte chunk per TMA box, no swizzle
int encode_sf_map(CUtensorMap* out, const void* base, long long chunks) {
    auto fn = ccmpi::drv<EncodeTiledFn>("cuTensorMapEncodeTiled");
    if (!fn) return CCMPI_ERR_CUDA;
    CCMPI_REQUIRE(base != nullptr && (reinterpret_cast<uintptr_t>(base) & 15) == 0, "scale buffer %p must be non-null and 16-byte aligned", base);
    CCMPI_CUDA_CHECK(cudaFree(nullptr));    // a current context (the buffer may live in the VMM-mapped heap)
    const cuuint64_t dims[2] = {128, (cuuint64_t)chunks};
    const cuuint64_t strides[1] = {512};
    const cuuint32_t box[2] = {128, 1};
    const cuuint32_t ones[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, ones,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        ccmpi::set_error("cuTensorMapEncodeTiled (scales) failed: CUresult %d", (int)r);
        return CCMPI_ERR_CUDA;
    }
    return CCMPI_OK;
}

constexpr int MX_BN = 128;

template <typename OutT, typename Policy>
int launch_mx(const Launch& lp, int grid, cudaStream_t stream, bool cooperative) {
    auto kern = gemm_kernel_2cta<MX_BN, false, false, OutT, Policy, 1, true>;
    constexpr int SMEM = Cfg2<MX_BN, Policy::kNeedsStaging, true>::SMEM_BYTES;
    static bool configured[64] = {};
    int dev_ = 0;
    CCMPI_CUDA_CHECK(cudaGetDevice(&dev_));
    if (!configured[dev_ & 63]) {
        CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        configured[dev_ & 63] = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = cooperative ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CCMPI_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, lp));
    return CCMPI_OK;
}

template <typename Policy>
int dispatch_mx(const Launch& lp, int out_dtype, int grid, cudaStream_t stream, bool cooperative) {
    if (out_dtype == CCMPI_BF16) return launch_mx<__nv_bfloat16, Policy>(lp, grid, stream, cooperative);
    return launch_mx<float, Policy>(lp, grid, stream, cooperative);
}

int check_mx(int M, int N, int K, long long lda, long long ldb, long long ldc, int out_dtype) {
    CCMPI_REQUIRE(M > 0 && N > 0 && K > 0 && M % 8 == 0 && N % 8 == 0 && K % 128 == 0,
                  "mxfp8 GEMM extents: M, N multiples of 8 and K a multiple of 128 (M=%d N=%d K=%d)", M, N, K);
    CCMPI_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && ldc % 8 == 0, "mxfp8 GEMM leading dimensions must keep rows 16-byte aligned");
    CCMPI_REQUIRE(out_dtype == CCMPI_BF16 || out_dtype == CCMPI_F32, "mxfp8 GEMM output must be bf16 or fp32");
    CCMPI_REQUIRE(use_2cta(), "block-scaled GEMMs exist in the CTA-pair kernel only (unset CCMPI_GEMM_1CTA)");
    return CCMPI_OK;
}

// operands of one rank: A [M][K] (+ optional second source A2 [a2_rows][a2_cols]), B [N][K], their scale arrays
int fill_mx(RankParams& rp, const void* A, const void* sfa, long long lda, const void* A2, const void* sfa2, long long lda2,
            int a2_rows, int a2_cols, const void* B, const void* sfb, long long ldb, int M, int N, int K) {
    int rc = ccmpi::make_tmap_2d(&rp.tmap_a, A, 1, M, K, lda, BM, 128);
    if (rc) return rc;
    rc = ccmpi::make_tmap_2d(&rp.tmap_a2, A2, 1, a2_rows, a2_cols, lda2, BM, 128);
    if (rc) return rc;
    rc = ccmpi::make_tmap_2d(&rp.tmap_b, B, 1, N, K, ldb, MX_BN / 2, 128);
    if (rc) return rc;
    const long long kc = K / 128, m_chunks = (M + 127) / 128, n_chunks = (N + 127) / 128;
    rc = encode_sf_map(&rp.tmap_sfa, sfa, m_chunks * kc);
    if (rc) return rc;
    rp.sf_a2_kblocks = a2_cols / 128;
    rc = encode_sf_map(&rp.tmap_sfa2, sfa2, (long long)((a2_rows + 127) / 128) * rp.sf_a2_kblocks);
    if (rc) return rc;
    return encode_sf_map(&rp.tmap_sfb, sfb, n_chunks * kc);
}

}  // namespace

extern "C" int ccmpi_gemm_mx(const void* A, const void* sfa, const void* B, const void* sfb, void* C, int M, int N, int K,
                             int lda, int ldb, int ldc, int out_dtype, int num_sms, void* stream) {
    int rc = check_mx(M, N, K, lda, ldb, ldc, out_dtype);
    if (rc) return rc;
    CCMPI_REQUIRE(A && B && C && sfa && sfb, "null operand");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count()) & ~1;
    CCMPI_REQUIRE(sms >= 2, "need at least one SM pair");
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = 1;
    RankParams& rp = lp.r[0];
    rc = fill_mx(rp, A, sfa, lda, A, sfa, lda, M, K, B, sfb, ldb, M, N, K);
    if (rc) return rc;
    rp.C = C;
    rp.ldc = ldc;
    rp.M = M; rp.N = N; rp.K = K;
    const int pair_tiles = (((M + BM - 1) / BM + 1) / 2) * ((N + MX_BN - 1) / MX_BN);
    const int npairs = pair_tiles < sms / 2 ? pair_tiles : sms / 2;
    return dispatch_mx<PlainPolicy>(lp, out_dtype, 2 * npairs, (cudaStream_t)stream, false);
}

// all-gather -> GEMM on the gathered K axis with block-scaled e4m3 operands.
//   xq_off    symmetric: this rank's shard            [M][K/n] e4m3 (row stride ldx bytes)
//   sfl_off   symmetric: its scales, chunked           [ceil(M/128)][K/n/128] x 512 B
//   xfull_off symmetric: gathered staging              [M][K] e4m3
//   sffull_off symmetric: gathered scales, chunked     [ceil(M/128)][K/128] x 512 B
//   B[v] / sfb[v]: this rank's weight shard [N][K] e4m3 and its chunked scales.
// Flags as in ccmpi_fused_ag_gemm (flag_target = running sum of ccmpi_fused_ag_ctas(2, ...), or 0).
extern "C" int ccmpi_fused_ag_gemm_mx(const ccmpi_dev_comm* comms, int nvr, size_t xq_off, int ldx, size_t sfl_off,
                                      size_t xfull_off, size_t sffull_off, const void* const* B, const void* const* sfb,
                                      void* const* C, int M, int N, int K, int ldb, int ldc, int out_dtype, int flag_base,
                                      unsigned epoch, unsigned flag_target, void* trace, int num_sms, void* stream) {
    int rc = check_mx(M, N, K, ldx, ldb, ldc, out_dtype);
    if (rc) return rc;
    CCMPI_REQUIRE(comms && nvr >= 1 && nvr <= CCMPI_MAX_VRANKS, "nvr=%d out of range", nvr);
    const int n = comms[0].size;
    CCMPI_REQUIRE(nvr == 1 || nvr == n, "emulation needs exactly one virtual rank per member (nvr=%d, size=%d)", nvr, n);
    const int m_blocks = (M + BM - 1) / BM;
    CCMPI_REQUIRE(flag_base >= 0 && flag_base + n * m_blocks <= CCMPI_PAD_USER_WORDS,
                  "fused op needs %d flag words at base %d, the pad has %d", n * m_blocks, flag_base, CCMPI_PAD_USER_WORDS);
    CCMPI_REQUIRE(K % n == 0 && (K / n) % 128 == 0, "K/n = %d must be a multiple of 128", K / n);
    CCMPI_REQUIRE(((xq_off | sfl_off | xfull_off | sffull_off) & 15) == 0, "symmetric offsets must be 16-byte aligned");
    const int chunk = K / n;        // fp8 elements = bytes
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const int budget = (sms / nvr) & ~1;
    CCMPI_REQUIRE(budget >= 2, "not enough SMs for %d emulated ranks", nvr);
    const int tiles = 2 * (((m_blocks + 1) / 2) * ((N + MX_BN - 1) / MX_BN));
    const int ctas = budget < tiles ? budget : tiles;
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        const ccmpi_dev_comm& c = comms[v];
        rc = fill_mx(rp, c.heap[c.rank] + xfull_off, c.heap[c.rank] + sffull_off, K, c.heap[c.rank] + xq_off,
                     c.heap[c.rank] + sfl_off, ldx, M, chunk, B[v], sfb[v], ldb, M, N, K);
        if (rc) return rc;
        rp.C = C[v];
        rp.ldc = ldc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = c;
        rp.off0 = xq_off; rp.off1 = xfull_off;
        rp.chunk = chunk / 2;               // the policies count K in 2-byte units (64 per K-block)
        rp.aux0 = ldx / 2;
        rp.flag_base = flag_base;
        rp.epoch = flag_target ? flag_target : epoch * (unsigned)ctas;
        rp.aux1 = 8;                        // bulk-copy push (scales travel with it), no copy into this rank's own staging
        rp.aux2 = 1;
        rp.sf_run_bytes = (chunk / 128) * 512;
        rp.off_sf_local = sfl_off;
        rp.off_sf_full = sffull_off;
        rp.trace = (unsigned long long*)trace;
    }
    return dispatch_mx<AgKPolicy>(lp, out_dtype, ctas * nvr, (cudaStream_t)stream, nvr > 1);
}

// GEMM -> reduce-scatter on N with block-scaled e4m3 operands (partials and result in `out_dtype`).
extern "C" int ccmpi_fused_gemm_rs_mx(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* sfa,
                                      const void* const* B, const void* const* sfb, size_t slots_off, void* const* C, int M,
                                      int N, int K, int lda, int ldb, int ldc, int out_dtype, int flag_base, unsigned epoch,
                                      const void* const* residual, long long res_ld, void* trace, int num_sms, void* stream) {
    int rc = check_mx(M, N, K, lda, ldb, ldc, out_dtype);
    if (rc) return rc;
    CCMPI_REQUIRE(comms && C && nvr >= 1 && nvr <= CCMPI_MAX_VRANKS, "nvr=%d out of range / null output", nvr);
    const int n = comms[0].size;
    CCMPI_REQUIRE(nvr == 1 || nvr == n, "emulation needs exactly one virtual rank per member (nvr=%d, size=%d)", nvr, n);
    CCMPI_REQUIRE(N % n == 0 && (N / n) % MX_BN == 0, "N/n = %d must be a multiple of %d", N / n, MX_BN);
    const int chunk = N / n;
    const int m_blocks = (M + BM - 1) / BM;
    const int flags = m_blocks * (chunk / MX_BN) * n;
    CCMPI_REQUIRE(flag_base >= 0 && flag_base + flags <= CCMPI_PAD_USER_WORDS,
                  "fused op needs %d flag words at base %d, the pad has %d", flags, flag_base, CCMPI_PAD_USER_WORDS);
    CCMPI_REQUIRE((slots_off & 15) == 0, "symmetric offsets must be 16-byte aligned");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const int tiles = 2 * (((m_blocks + 1) / 2) * (N / MX_BN));
    int per_rank = (sms / nvr) & ~1;
    if (per_rank > tiles) per_rank = tiles;
    CCMPI_REQUIRE(per_rank >= 2, "not enough SMs for %d emulated ranks", nvr);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        rc = fill_mx(rp, A[v], sfa[v], lda, A[v], sfa[v], lda, M, K, B[v], sfb[v], ldb, M, N, K);
        if (rc) return rc;
        rp.C = C[v];
        rp.ldc = ldc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = comms[v];
        rp.off1 = slots_off;
        rp.chunk = chunk;
        rp.flag_base = flag_base;
        rp.epoch = epoch;
        rp.aux0 = MX_BN;
        if (residual) {
            CCMPI_REQUIRE(out_dtype == CCMPI_BF16 && residual[v] != nullptr && res_ld % 8 == 0 && res_ld >= chunk,
                          "reduce-scatter residual: bf16 output, non-null pointer, row stride %% 8 == 0 and >= N/n");
            rp.epi_op = EPI_ADD;
            rp.epi_in = residual[v];
            rp.epi_ld = res_ld;
        }
        rp.trace = (unsigned long long*)trace;
    }
    return dispatch_mx<RsNPolicy>(lp, out_dtype, per_rank * nvr, (cudaStream_t)stream, nvr > 1);
}

// ------------------------------------------------------------------------------------------------
// Quantiser: bf16 [R][K] -> e4m3 [R][K] and one UE8M0 scale per 1 x 32 block, written straight into the chunked
// layout the tensor cores read (512-byte chunks of 128 rows x 4 scales, K-fastest; inside a chunk
// (r % 32) * 16 + (r / 32) * 4 + k). One thread per block of 32 elements: 64 bytes in, 32 + 1 bytes out.
// scale = 2^ceil(log2(amax / 448)), so the block's largest magnitude fits e4m3's range.
// ------------------------------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) quantize_mxfp8_kernel(const __nv_bfloat16* __restrict__ x, long long ldx,
                                                             uint8_t* __restrict__ q, long long ldq, uint8_t* __restrict__ sf,
                                                             int R, int K) {
    const int bpr = K / 32;                                   // blocks per row
    const long long total = (long long)R * bpr;
    const int kchunks = (K + 127) / 128;
    for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int r = (int)(i / bpr), b = (int)(i - (long long)r * bpr);
        const uint4* src = reinterpret_cast<const uint4*>(x + (long long)r * ldx + b * 32);
        float v[32];
        float amax = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint4 u = src[j];
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __bfloat1622float2(h[e]);
                v[8 * j + 2 * e] = f.x;
                v[8 * j + 2 * e + 1] = f.y;
                amax = fmaxf(amax, fmaxf(fabsf(f.x), fabsf(f.y)));
            }
        }
        amax = fmaxf(amax, 1.1754944e-38f);                   // 2^-126
        int e = (int)ceilf(log2f(__fdiv_rn(amax, 448.0f)));
        e = max(-127, min(127, e));
        const float inv = exp2f((float)-e);
        uint32_t out[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint32_t w = 0;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float s = fminf(fmaxf(v[4 * j + t] * inv, -448.0f), 448.0f);
                w |= (uint32_t)__nv_cvt_float_to_fp8(s, __NV_SATFINITE, __NV_E4M3) << (8 * t);
            }
            out[j] = w;
        }
        uint4* dst = reinterpret_cast<uint4*>(q + (long long)r * ldq + b * 32);
        dst[0] = make_uint4(out[0], out[1], out[2], out[3]);
        dst[1] = make_uint4(out[4], out[5], out[6], out[7]);
        const int chunk = (r >> 7) * kchunks + (b >> 2);
        const int rr = r & 127;
        sf[(long long)chunk * 512 + (rr & 31) * 16 + (rr >> 5) * 4 + (b & 3)] = (uint8_t)(e + 127);
    }
}
}  // namespace

// sf must hold ceil(R/128) * ceil(K/128) * 512 bytes; rows beyond R and scale columns beyond K/32 of the last
// chunks are left untouched (fill the buffer with the neutral exponent 127 once if R or K/32 is ragged).
extern "C" int ccmpi_quantize_mxfp8(const void* x, long long ldx, void* q, long long ldq, void* sf, int R, int K, void* stream) {
    CCMPI_REQUIRE(x && q && sf && R > 0 && K > 0 && K % 32 == 0 && ldx % 8 == 0 && ldq % 16 == 0,
                  "quantize_mxfp8: K %% 32 == 0, 16-byte aligned rows (R=%d K=%d)", R, K);
    CCMPI_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q)) & 15) == 0, "quantize_mxfp8: 16-byte aligned pointers");
    const long long total = (long long)R * (K / 32);
    long long grid = (total + 255) / 256;
    const long long cap = (long long)ccmpi::device_sm_count() * 8;
    if (grid > cap) grid = cap;
    quantize_mxfp8_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ldx, (uint8_t*)q, ldq, (uint8_t*)sf, R, K);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}
