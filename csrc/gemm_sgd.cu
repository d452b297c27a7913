// Weight-gradient GEMM with the SGD update in its epilogue:  W[M][N] -= lr * (A * B).
//
// For a layer whose gradient needs no data-parallel reduction (dp = 1) the usual sequence is
//   dW = x^T dy   (write M*N)      W -= lr * dW   (read M*N twice, write M*N)
// Here the tile's fp32 accumulator never leaves the SM as a gradient: the epilogue reads the W tile,
// subtracts lr * accumulator in fp32 and writes W back -- one read and one write of W in total, and
// the gradient is applied unrounded. The reference has no optimizer (or GEMM); this belongs to the
// training step the new framework adds around its communicators.
//
// The new weight is rounded to bf16 STOCHASTICALLY (sr.cuh; aux1 = step counter, aux2 = seed; aux1 < 0
// selects round-to-nearest): a step far below one bf16 ulp still moves the weights in expectation, with
// no fp32 master copy. CTA-pair kernel, bf16 weights; the models use it when dp = 1.
#include "gemm_host.h"
#include "sr.cuh"

namespace gemm {

struct SgdPolicy : PlainPolicy {
    static constexpr bool kEpiOps = false;   // own store: no fused elementwise epilogue
    template <typename OutT>
    __device__ static void store(const RankParams& p, const TileCoord& tc, int row, int col0, const uint32_t (&v)[32]) {
        static_assert(sizeof(OutT) == 2, "weights are bf16");
        if (row >= p.M) return;
        const int ncols = min(32, p.N - col0);
        if (ncols <= 0) return;
        const float lr = __int_as_float(p.aux0);
        const bool stochastic = p.aux1 >= 0;
        const unsigned long long pair0 = ((unsigned long long)row * (unsigned long long)p.ldc + (unsigned long long)col0) >> 1;
        __nv_bfloat16* w = reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (j * 8 < ncols) {
                const uint4 old = ptx::ld_v4(w + 8 * j);
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&old);
                uint4 q;
                uint32_t* qo = reinterpret_cast<uint32_t*>(&q);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 f = __bfloat1622float2(h[e]);
                    const float lo = sr::axpy(f.x, lr, __uint_as_float(v[8 * j + 2 * e]));
                    const float hi = sr::axpy(f.y, lr, __uint_as_float(v[8 * j + 2 * e + 1]));
                    qo[e] = stochastic ? sr::pack2(lo, hi, sr::bits(pair0 + 4 * j + e, (uint32_t)p.aux1, (uint32_t)p.aux2))
                                       : pack_bf16(__float_as_uint(lo), __float_as_uint(hi));
                }
                ptx::st_v4(w + 8 * j, q);
            }
        }
    }
};

template <int BN, bool A_MN, bool B_MN>
static int launch_sgd(const Launch& lp, int grid, cudaStream_t stream) {
    auto kern = gemm_kernel_2cta<BN, A_MN, B_MN, __nv_bfloat16, SgdPolicy>;
    static bool configured[64] = {};   // the attribute is per device (and per instantiation)
    int dev_ = 0;
    CCMPI_CUDA_CHECK(cudaGetDevice(&dev_));
    if (!configured[dev_ & 63]) {
        CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2<BN, false>::SMEM_BYTES));
        configured[dev_ & 63] = true;
    }
    kern<<<grid, NUM_THREADS, Cfg2<BN, false>::SMEM_BYTES, stream>>>(lp);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

}  // namespace gemm

extern "C" int ccmpi_gemm_sgd_bf16(const void* A, const void* B, void* W, int M, int N, int K, int lda, int ldb, int ldw,
                                   int a_mn_major, int b_mn_major, float lr, int num_sms, void* stream) {
    return ccmpi_gemm_sgd_bf16_sr(A, B, W, M, N, K, lda, ldb, ldw, a_mn_major, b_mn_major, lr, -1, 0, num_sms, stream);
}

// `step` >= 0: stochastic rounding of the updated weights, random bits from (element, step, seed);
// `step` < 0: round to nearest.
extern "C" int ccmpi_gemm_sgd_bf16_sr(const void* A, const void* B, void* W, int M, int N, int K, int lda, int ldb, int ldw,
                                      int a_mn_major, int b_mn_major, float lr, int step, unsigned seed, int num_sms,
                                      void* stream) {
    using namespace gemm;
    int rc = check_shapes(M, N, K, lda, ldb, ldw, CCMPI_BF16);
    if (rc) return rc;
    CCMPI_REQUIRE(W != nullptr, "null weight pointer");
    CCMPI_REQUIRE(use_2cta(), "the GEMM + SGD kernel exists for the CTA-pair GEMM only (unset CCMPI_GEMM_1CTA)");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count()) & ~1;
    CCMPI_REQUIRE(sms >= 2, "need at least one SM pair");
    const int BN = pick_bn(M, N, sms);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = 1;
    RankParams& rp = lp.r[0];
    rc = fill_tmaps(rp, A, A, B, M, N, K, lda, lda, ldb, a_mn_major != 0, b_mn_major != 0, BN / 2, a_mn_major ? K : M,
                    a_mn_major ? M : K);
    if (rc) return rc;
    rp.C = W;
    rp.ldc = ldw;
    rp.M = M; rp.N = N; rp.K = K;
    rp.aux0 = __builtin_bit_cast(int, lr);
    rp.aux1 = step;
    rp.aux2 = (int)seed;
    const int pair_tiles = (((M + BM - 1) / BM + 1) / 2) * ((N + BN - 1) / BN);
    const int npairs = pair_tiles < sms / 2 ? pair_tiles : sms / 2;
    const cudaStream_t st = (cudaStream_t)stream;
#define CCMPI_SGD_CASE(bn, amn, bmn) \
    if (BN == bn && (a_mn_major != 0) == amn && (b_mn_major != 0) == bmn) return launch_sgd<bn, amn, bmn>(lp, 2 * npairs, st);
    CCMPI_SGD_CASE(256, false, false)
    CCMPI_SGD_CASE(256, false, true)
    CCMPI_SGD_CASE(256, true, false)
    CCMPI_SGD_CASE(256, true, true)
    CCMPI_SGD_CASE(128, false, false)
    CCMPI_SGD_CASE(128, false, true)
    CCMPI_SGD_CASE(128, true, false)
    CCMPI_SGD_CASE(128, true, true)
#undef CCMPI_SGD_CASE
    ccmpi::set_error("no CTA-pair GEMM instantiation for BN=%d", BN);
    return CCMPI_ERR_UNSUPPORTED;
}
