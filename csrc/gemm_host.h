// Host-side helpers shared by gemm_tcgen05.cu (plain GEMM) and fused_gemm.cu.
#pragma once
#include "gemm_kernel.cuh"
#include "host_common.h"

namespace gemm {

int fill_tmaps(RankParams& rp, const void* A, const void* A2, const void* B, int M, int N, int K, long long lda,
               long long lda2, long long ldb, bool a_mn, bool b_mn, int BN, int a2_rows, int a2_cols);
int pick_bn(int M, int N, int sms);
int check_shapes(int M, int N, int K, long long lda, long long ldb, long long ldc, int out_dtype);

template <int BN, bool A_MN, bool B_MN, typename OutT, typename Policy>
int launch(const Launch& lp, int grid, cudaStream_t stream, bool cooperative) {
    auto kern = gemm_kernel<BN, A_MN, B_MN, OutT, Policy>;
    static bool configured = false;  // per instantiation
    if (!configured) {
        CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
        configured = true;
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
