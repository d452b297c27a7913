// fp8 (e4m3) GEMM on the CTA-pair tcgen05 kernel: C[M][N] = scale * A[M][K] * B[N][K]^T, both
// operands K-major, fp32 accumulation in TMEM, bf16 or fp32 output. `scale` is the product of the
// two per-tensor dequantisation scales and is applied in the epilogue.
//
// Run on a B200 in round 2 (tests/test_gpu_fp8.py; UTCQMMA.2CTA in the SASS). Nothing in the
// default paths calls it: it is the per-tensor-scaled step towards BASELINE.md config 3, whose
// block-scaled form (tcgen05.mma.kind::mxf8f6f4.block_scale, scale-factor tiles in tensor memory)
// is csrc/gemm_mxfp8.cu (stand-alone) and the MX mode of the generic CTA-pair kernel
// (csrc/fused_gemm_mx.cu: plain, all-gather->GEMM, GEMM->reduce-scatter).
#include "gemm_host.h"

namespace gemm {

struct ScaledPolicy : PlainPolicy {
    static constexpr bool kEpiOps = false;   // own store: no fused elementwise epilogue
    template <typename OutT>
    __device__ static void store(const RankParams& p, const TileCoord& tc, int row, int col0, const uint32_t (&v)[32]) {
        if (row >= p.M) return;
        const int ncols = min(32, p.N - col0);
        if (ncols <= 0) return;
        const float s = __int_as_float(p.aux0);
        uint32_t w[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) w[i] = __float_as_uint(__uint_as_float(v[i]) * s);
        store_row32<OutT>(reinterpret_cast<OutT*>(p.C) + (long long)row * p.ldc + col0, w, ncols, false);
    }
};

template <int BN, typename OutT>
static int launch_fp8(const Launch& lp, int grid, cudaStream_t stream) {
    auto kern = gemm_kernel_2cta<BN, false, false, OutT, ScaledPolicy, 1>;
    CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2<BN, false>::SMEM_BYTES));
    kern<<<grid, NUM_THREADS, Cfg2<BN, false>::SMEM_BYTES, stream>>>(lp);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

}  // namespace gemm

extern "C" int ccmpi_gemm_fp8_e4m3(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                                   int ldc, float scale, int out_dtype, int num_sms, void* stream) {
    using namespace gemm;
    CCMPI_REQUIRE(A && B && C, "null operand");
    CCMPI_REQUIRE(M > 0 && N > 0 && K > 0 && M % 8 == 0 && N % 8 == 0 && K % 16 == 0,
                  "fp8 GEMM extents: M, N multiples of 8 and K a multiple of 16 (M=%d N=%d K=%d)", M, N, K);
    CCMPI_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && ldc % 8 == 0, "fp8 GEMM leading dimensions must keep rows 16-byte aligned");
    CCMPI_REQUIRE(out_dtype == CCMPI_BF16 || out_dtype == CCMPI_F32, "fp8 GEMM output must be bf16 or fp32");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count()) & ~1;
    CCMPI_REQUIRE(sms >= 2, "need at least one SM pair");
    const int BN = pick_bn(M, N, sms);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = 1;
    RankParams& rp = lp.r[0];
    int rc = ccmpi::make_tmap_2d(&rp.tmap_a, A, 1, M, K, lda, BM, 128);
    if (rc) return rc;
    rp.tmap_a2 = rp.tmap_a;
    rc = ccmpi::make_tmap_2d(&rp.tmap_b, B, 1, N, K, ldb, BN / 2, 128);
    if (rc) return rc;
    rp.C = C;
    rp.ldc = ldc;
    rp.M = M; rp.N = N; rp.K = K;
    memcpy(&rp.aux0, &scale, sizeof(float));
    const int pair_tiles = (((M + BM - 1) / BM + 1) / 2) * ((N + BN - 1) / BN);
    const int npairs = pair_tiles < sms / 2 ? pair_tiles : sms / 2;
    cudaStream_t st = (cudaStream_t)stream;
    if (BN == 256) return out_dtype == CCMPI_BF16 ? launch_fp8<256, __nv_bfloat16>(lp, 2 * npairs, st) : launch_fp8<256, float>(lp, 2 * npairs, st);
    return out_dtype == CCMPI_BF16 ? launch_fp8<128, __nv_bfloat16>(lp, 2 * npairs, st) : launch_fp8<128, float>(lp, 2 * npairs, st);
}
