// Host-side helpers shared by gemm_tcgen05.cu (plain GEMM) and fused_gemm.cu.
#pragma once
#include "gemm_kernel_2cta.cuh"
#include "host_common.h"

namespace gemm {

int fill_tmaps(RankParams& rp, const void* A, const void* A2, const void* B, int M, int N, int K, long long lda,
               long long lda2, long long ldb, bool a_mn, bool b_mn, int b_box_rows, int a2_rows, int a2_cols);
int pick_bn(int M, int N, int sms);
// CTA-pair kernels are the default; CCMPI_GEMM_1CTA=1 selects the single-CTA kernels
bool use_2cta();
int check_shapes(int M, int N, int K, long long lda, long long ldb, long long ldc, int out_dtype);
int set_epilogue(RankParams& rp, int epi_op, const void* epi_in, void* epi_out2, long long epi_ld, int N, int out_dtype,
                 bool pairs);

template <int BN, bool A_MN, bool B_MN, typename OutT, typename Policy>
int launch(const Launch& lp, int grid, cudaStream_t stream, bool cooperative) {
    auto kern = gemm_kernel<BN, A_MN, B_MN, OutT, Policy>;
    static bool configured[64] = {};   // the attribute is per device (and per instantiation)
    int dev_ = 0;
    CCMPI_CUDA_CHECK(cudaGetDevice(&dev_));
    if (!configured[dev_ & 63]) {
        CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
        configured[dev_ & 63] = true;
    }
    if (!cooperative) {
        kern<<<grid, NUM_THREADS, Cfg<BN>::SMEM_BYTES, stream>>>(lp);
    } else {
        // CTAs of different (virtual) ranks wait on one another: co-residency must be guaranteed
        void* args[] = {const_cast<Launch*>(&lp)};
        CCMPI_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(NUM_THREADS), args,
                                                     Cfg<BN>::SMEM_BYTES, stream));
    }
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

template <int BN, bool A_MN, bool B_MN, typename OutT, typename Policy>
int launch2(const Launch& lp, int grid, cudaStream_t stream, bool cooperative) {
    auto kern = gemm_kernel_2cta<BN, A_MN, B_MN, OutT, Policy>;
    static bool configured[64] = {};   // the attribute is per device (and per instantiation)
    int dev_ = 0;
    CCMPI_CUDA_CHECK(cudaGetDevice(&dev_));
    if (!configured[dev_ & 63]) {
        CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2<BN, Policy::kNeedsStaging>::SMEM_BYTES));
        configured[dev_ & 63] = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid);          // even: the kernel carries __cluster_dims__(2,1,1)
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = Cfg2<BN, Policy::kNeedsStaging>::SMEM_BYTES;
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
int dispatch2(const Launch& lp, int BN, bool a_mn, bool b_mn, int out_dtype, int grid, cudaStream_t stream,
              bool cooperative) {
#define CCMPI_GEMM_CASE2(bn, amn, bmn)                                                                  \
    if (BN == bn && a_mn == amn && b_mn == bmn) {                                                       \
        if (out_dtype == CCMPI_BF16) return launch2<bn, amn, bmn, __nv_bfloat16, Policy>(lp, grid, stream, cooperative); \
        return launch2<bn, amn, bmn, float, Policy>(lp, grid, stream, cooperative);                    \
    }
    CCMPI_GEMM_CASE2(256, false, false)
    CCMPI_GEMM_CASE2(256, false, true)
    CCMPI_GEMM_CASE2(256, true, false)
    CCMPI_GEMM_CASE2(256, true, true)
    CCMPI_GEMM_CASE2(128, false, false)
    CCMPI_GEMM_CASE2(128, false, true)
    CCMPI_GEMM_CASE2(128, true, false)
    CCMPI_GEMM_CASE2(128, true, true)
#undef CCMPI_GEMM_CASE2
    ccmpi::set_error("no CTA-pair GEMM instantiation for BN=%d", BN);
    return CCMPI_ERR_UNSUPPORTED;
}

template <typename Policy>
int dispatch(const Launch& lp, int BN, bool a_mn, bool b_mn, int out_dtype, int grid, cudaStream_t stream,
             bool cooperative) {
#define CCMPI_GEMM_CASE(bn, amn, bmn)                                                                   \
    if (BN == bn && a_mn == amn && b_mn == bmn) {                                                       \
        if (out_dtype == CCMPI_BF16) return launch<bn, amn, bmn, __nv_bfloat16, Policy>(lp, grid, stream, cooperative); \
        return launch<bn, amn, bmn, float, Policy>(lp, grid, stream, cooperative);                     \
    }
    CCMPI_GEMM_CASE(256, false, false)
    CCMPI_GEMM_CASE(256, false, true)
    CCMPI_GEMM_CASE(256, true, false)
    CCMPI_GEMM_CASE(256, true, true)
    CCMPI_GEMM_CASE(128, false, false)
    CCMPI_GEMM_CASE(128, false, true)
    CCMPI_GEMM_CASE(128, true, false)
    CCMPI_GEMM_CASE(128, true, true)
#undef CCMPI_GEMM_CASE
    ccmpi::set_error("no GEMM instantiation for BN=%d", BN);
    return CCMPI_ERR_UNSUPPORTED;
}

}  // namespace gemm
