// Plain bf16 GEMM entry point on the tcgen05 mainloop. The reference ships no GEMM at all
// (README.md:173-175 says the FC compute is "taken care of" elsewhere); these are the three
// products a tensor-parallel FC layer needs around func_impl.py:76-187:
//     y = x W        (A K-major,  B stored [K][N] -> b_mn_major)
//     dx = dy W^T    (A K-major,  B stored [N][K] -> K-major)
//     dW = x^T dy    (A stored [K][M] -> a_mn_major, B stored [K][N] -> b_mn_major)
#include <cstdlib>

#include "gemm_host.h"

namespace gemm {

// Fill the three TMA descriptors of a rank. A2 (alternate A source) defaults to A.
int fill_tmaps(RankParams& rp, const void* A, const void* A2, const void* B, int M, int N, int K, long long lda,
               long long lda2, long long ldb, bool a_mn, bool b_mn, int b_box_rows, int a2_rows, int a2_cols) {
    int rc;
    if (!a_mn) {
        rc = ccmpi::make_tmap_2d(&rp.tmap_a, A, 2, M, K, lda, BM, BK);
        if (rc) return rc;
        rc = ccmpi::make_tmap_2d(&rp.tmap_a2, A2, 2, a2_rows, a2_cols, lda2, BM, BK);
    } else {
        rc = ccmpi::make_tmap_2d(&rp.tmap_a, A, 2, K, M, lda, BK, 64);
        if (rc) return rc;
        rc = ccmpi::make_tmap_2d(&rp.tmap_a2, A2, 2, a2_rows, a2_cols, lda2, BK, 64);
    }
    if (rc) return rc;
    if (!b_mn)
        rc = ccmpi::make_tmap_2d(&rp.tmap_b, B, 2, N, K, ldb, b_box_rows, BK);
    else
        rc = ccmpi::make_tmap_2d(&rp.tmap_b, B, 2, K, N, ldb, BK, 64);
    return rc;
}

int pick_bn(int M, int N, int sms) {
    auto eff = [&](int bn) {
        const long long tiles = (long long)((M + BM - 1) / BM) * ((N + bn - 1) / bn);
        const long long waves = (tiles + sms - 1) / sms;
        // useful fraction of issued tile slots, discounted by edge waste in N
        const double edge = (double)N / (double)(((N + bn - 1) / bn) * bn);
        return edge * (double)tiles / (double)(waves * sms);
    };
    if (N <= 128) return 128;
    // the 256-wide tile does ~1.4x the work per byte moved; the narrow one only pays off when the
    // wide one would leave a large part of the last wave idle
    return (eff(256) >= 0.75 * eff(128)) ? 256 : 128;
}

bool use_2cta() {
    static const bool v = [] {
        const char* e = getenv("CCMPI_GEMM_1CTA");
        return !(e && e[0] == '1');
    }();
    return v;
}

int check_shapes(int M, int N, int K, long long lda, long long ldb, long long ldc, int out_dtype) {
    CCMPI_REQUIRE(M > 0 && N > 0 && K > 0, "GEMM extents must be positive (M=%d N=%d K=%d)", M, N, K);
    CCMPI_REQUIRE(K % 8 == 0 && N % 8 == 0 && M % 8 == 0, "GEMM extents must be multiples of 8 (M=%d N=%d K=%d)", M, N, K);
    CCMPI_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "GEMM leading dimensions must be multiples of 8");
    CCMPI_REQUIRE(out_dtype == CCMPI_BF16 || out_dtype == CCMPI_F32, "GEMM output must be bf16 or fp32");
    return CCMPI_OK;
}

}  // namespace gemm

namespace gemm {
// Checks shared by every entry point that takes a fused elementwise epilogue; fills the RankParams fields.
int set_epilogue(RankParams& rp, int epi_op, const void* epi_in, void* epi_out2, long long epi_ld, int N, int out_dtype,
                 bool pairs) {
    rp.epi_op = EPI_NONE;
    if (epi_op == EPI_NONE) return CCMPI_OK;
    CCMPI_REQUIRE(epi_op >= EPI_GELU_DUAL && epi_op <= EPI_ADD, "unknown epilogue op %d", epi_op);
    CCMPI_REQUIRE(pairs, "fused epilogues exist in the CTA-pair GEMM only (unset CCMPI_GEMM_1CTA)");
    CCMPI_REQUIRE(out_dtype == CCMPI_BF16, "fused epilogues need a bf16 output");
    CCMPI_REQUIRE(N % 32 == 0, "fused epilogues need N %% 32 == 0 (N = %d)", N);
    CCMPI_REQUIRE(epi_ld % 8 == 0 && epi_ld >= N, "epilogue operand row stride %lld must be a multiple of 8 and >= N", epi_ld);
    if (epi_op == EPI_GELU_DUAL) CCMPI_REQUIRE(epi_out2 != nullptr, "EPI_GELU_DUAL needs a second output");
    else CCMPI_REQUIRE(epi_in != nullptr, "this epilogue needs a second operand");
    rp.epi_op = epi_op;
    rp.epi_in = epi_in;
    rp.epi_out2 = epi_out2;
    rp.epi_ld = epi_ld;
    return CCMPI_OK;
}
}  // namespace gemm

// C = A * B with optional elementwise work in the epilogue (EpiOp in gemm_tcgen05.cuh): 1 = also write
// gelu(C) to epi_out2, 2 = multiply by gelu'(epi_in), 3 = add epi_in. `trace`: TRACE_WORDS uint64 per CTA.
extern "C" int ccmpi_gemm_bf16_ex(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                                  int a_mn_major, int b_mn_major, int out_dtype, int accumulate, int epi_op,
                                  const void* epi_in, void* epi_out2, long long epi_ld, void* trace, int num_sms,
                                  void* stream) {
    using namespace gemm;
    int rc = check_shapes(M, N, K, lda, ldb, ldc, out_dtype);
    if (rc) return rc;
    CCMPI_REQUIRE(!accumulate || out_dtype == CCMPI_F32, "accumulate needs an fp32 output");
    const int sms = num_sms > 0 ? num_sms : ccmpi::device_sm_count();
    const bool pairs = use_2cta() && sms >= 2;
    const int BN = pick_bn(M, N, pairs ? sms / 2 * 2 : sms);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = 1;
    RankParams& rp = lp.r[0];
    rc = fill_tmaps(rp, A, A, B, M, N, K, lda, lda, ldb, a_mn_major != 0, b_mn_major != 0, pairs ? BN / 2 : BN,
                    a_mn_major ? K : M, a_mn_major ? M : K);
    if (rc) return rc;
    rp.C = C;
    rp.ldc = ldc;
    rp.M = M;
    rp.N = N;
    rp.K = K;
    rp.accumulate = accumulate;
    rc = set_epilogue(rp, epi_op, epi_in, epi_out2, epi_ld, N, out_dtype, pairs);
    if (rc) return rc;
    rp.trace = pairs ? (unsigned long long*)trace : nullptr;
    if (pairs) {
        const int pair_tiles = (((M + BM - 1) / BM + 1) / 2) * ((N + BN - 1) / BN);
        const int npairs = pair_tiles < sms / 2 ? pair_tiles : sms / 2;
        return dispatch2<PlainPolicy>(lp, BN, a_mn_major != 0, b_mn_major != 0, out_dtype, 2 * npairs, (cudaStream_t)stream, false);
    }
    const int tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    const int grid = tiles < sms ? tiles : sms;
    return dispatch<PlainPolicy>(lp, BN, a_mn_major != 0, b_mn_major != 0, out_dtype, grid, (cudaStream_t)stream, false);
}

extern "C" int ccmpi_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                               int a_mn_major, int b_mn_major, int out_dtype, int accumulate, int num_sms,
                               void* stream) {
    return ccmpi_gemm_bf16_ex(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_dtype, accumulate, 0, nullptr,
                              nullptr, 0, nullptr, num_sms, stream);
}
