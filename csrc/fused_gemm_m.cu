// Host entry points of the token-axis (sequence-parallel) fused kernels; device side:
// fused_policies_m.cuh. Run on B200s in round 2 (tests/test_gpu_seq.py).
//
// Same calling conventions as fused_gemm.cu: `nvr` device communicators (1 in production, one per
// member to emulate the ranks in one cooperative launch on one GPU), per-rank pointer lists, and
// symmetric-heap offsets for everything peers touch. CTA-pair kernels only.
#include "fused_policies_m.cuh"
#include "gemm_host.h"

using namespace gemm;

namespace {

int checks_m(const ccmpi_dev_comm* comms, int nvr, int M, int flag_words, int flag_base, int* rows_per_rank) {
    CCMPI_REQUIRE(comms && nvr >= 1 && nvr <= CCMPI_MAX_VRANKS, "nvr=%d out of range", nvr);
    const int n = comms[0].size;
    CCMPI_REQUIRE(nvr == 1 || nvr == n, "emulation needs exactly one virtual rank per member (nvr=%d, size=%d)", nvr, n);
    CCMPI_REQUIRE(use_2cta(), "the token-axis fused kernels exist for the CTA-pair GEMM only (unset CCMPI_GEMM_1CTA)");
    CCMPI_REQUIRE(M % n == 0 && (M / n) % (2 * BM) == 0,
                  "M/n = %d rows per rank must be a multiple of %d (a CTA pair never straddles two owners)", M / n, 2 * BM);
    CCMPI_REQUIRE(flag_base >= 0 && flag_base + flag_words <= CCMPI_PAD_USER_WORDS,
                  "fused op needs %d flag words at base %d, the pad has %d", flag_words, flag_base, CCMPI_PAD_USER_WORDS);
    *rows_per_rank = M / n;
    return CCMPI_OK;
}

// AgM never has an MN-major A (the gathered activation is row-major): half the instantiations
int dispatch_agm(const Launch& lp, int BN, bool b_mn, int out_dtype, int grid, cudaStream_t stream, bool coop) {
#define CCMPI_AGM_CASE(bn, bmn)                                                                              \
    if (BN == bn && b_mn == bmn) {                                                                           \
        if (out_dtype == CCMPI_BF16) return launch2<bn, false, bmn, __nv_bfloat16, AgMPolicy>(lp, grid, stream, coop); \
        return launch2<bn, false, bmn, float, AgMPolicy>(lp, grid, stream, coop);                           \
    }
    CCMPI_AGM_CASE(256, false)
    CCMPI_AGM_CASE(256, true)
    CCMPI_AGM_CASE(128, false)
    CCMPI_AGM_CASE(128, true)
#undef CCMPI_AGM_CASE
    ccmpi::set_error("no CTA-pair GEMM instantiation for BN=%d", BN);
    return CCMPI_ERR_UNSUPPORTED;
}

}  // namespace

// all-gather -> GEMM, gather on M:  C_r[M][N] = [x_0 ; x_1 ; ... ; x_{n-1}] * B_r,  x_r = rank r's [M/n][K] rows.
// x_off: symmetric [M/n][K] (row stride ldx), xfull_off: symmetric [M][K] that every member's rows are pushed into.
extern "C" int ccmpi_fused_ag_gemm_m(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                                     const void* const* B, void* const* C, int M, int N, int K, int ldb, int ldc,
                                     int b_kn, int out_dtype, int flag_base, unsigned epoch, unsigned flag_target,
                                     int num_sms, void* stream) {
    int rc = check_shapes(M, N, K, ldx, ldb, ldc, out_dtype);
    if (rc) return rc;
    int Mr = 0;
    rc = checks_m(comms, nvr, M, (M + BM - 1) / BM, flag_base, &Mr);
    if (rc) return rc;
    CCMPI_REQUIRE(K % BK == 0 && (K * 2) % 128 == 0, "K = %d must be a multiple of %d", K, BK);
    CCMPI_REQUIRE((x_off | xfull_off) % 16 == 0, "symmetric offsets must be 16-byte aligned");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const int budget = (sms / nvr) & ~1;
    CCMPI_REQUIRE(budget >= 2, "not enough SMs for %d emulated ranks", nvr);
    const int BN = pick_bn(M, N, budget);
    const int m_blocks = M / BM;
    const int tiles = 2 * ((m_blocks / 2) * ((N + BN - 1) / BN));
    const int ctas = budget < tiles ? budget : tiles;
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        const ccmpi_dev_comm& c = comms[v];
        rc = fill_tmaps(rp, c.heap[c.rank] + xfull_off, c.heap[c.rank] + x_off, B[v], M, N, K, K, ldx, ldb, false,
                        b_kn != 0, BN / 2, Mr, K);
        if (rc) return rc;
        rp.C = C[v];
        rp.ldc = ldc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = c;
        rp.off0 = x_off; rp.off1 = xfull_off;
        rp.chunk = Mr;
        rp.flag_base = flag_base;
        rp.epoch = flag_target ? flag_target : epoch * (unsigned)ctas;   // see ccmpi_fused_ag_gemm
        rp.aux0 = ldx;
        rp.aux2 = 1;
    }
    return dispatch_agm(lp, BN, b_kn != 0, out_dtype, ctas * nvr, (cudaStream_t)stream, nvr > 1);
}

// GEMM -> reduce-scatter, scatter on M:  C_r[M/n][N] = sum over ranks s of (A_s * B_s)[rows of rank r].
// slots_off: symmetric [n sources][M/n][N] of the output dtype.
extern "C" int ccmpi_fused_gemm_rs_m(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                                     size_t slots_off, void* const* C, int M, int N, int K, int lda, int ldb, int ldc,
                                     int a_km, int b_kn, int out_dtype, int flag_base, unsigned epoch, int num_sms,
                                     void* stream) {
    int rc = check_shapes(M, N, K, lda, ldb, ldc, out_dtype);
    if (rc) return rc;
    CCMPI_REQUIRE(C != nullptr, "gemm_rs_m needs an output pointer per rank");
    CCMPI_REQUIRE(N % 128 == 0, "N = %d must be a multiple of 128", N);
    const int BN = (N % 256 == 0) ? 256 : 128;
    const int n = comms ? comms[0].size : 1;
    int Mr = 0;
    rc = checks_m(comms, nvr, M, (M / n / BM) * (N / BN) * n, flag_base, &Mr);
    if (rc) return rc;
    CCMPI_REQUIRE(slots_off % 16 == 0, "symmetric offsets must be 16-byte aligned");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const int tiles = 2 * ((M / BM / 2) * (N / BN));
    int per_rank = (sms / nvr) & ~1;
    if (per_rank > tiles) per_rank = tiles;
    CCMPI_REQUIRE(per_rank >= 2, "not enough SMs for %d emulated ranks", nvr);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        rc = fill_tmaps(rp, A[v], A[v], B[v], M, N, K, lda, lda, ldb, a_km != 0, b_kn != 0, BN / 2, a_km ? K : M,
                        a_km ? M : K);
        if (rc) return rc;
        rp.C = C[v];
        rp.ldc = ldc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = comms[v];
        rp.off1 = slots_off;
        rp.chunk = Mr;
        rp.flag_base = flag_base;
        rp.epoch = epoch;
        rp.aux0 = BN;
    }
    return dispatch2<RsMPolicy>(lp, BN, a_km != 0, b_kn != 0, out_dtype, per_rank * nvr, (cudaStream_t)stream, nvr > 1);
}

// all-gather -> GEMM -> all-gather: y_full[M][N*n] (symmetric) = [x_0 | ... | x_{n-1}] * B_r in column block r,
// on every member; x_r = rank r's [M][K/n] shard (K-gather as in ccmpi_fused_ag_gemm).
extern "C" int ccmpi_fused_ag_gemm_ag(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                                      const void* const* B, size_t yfull_off, int M, int N, int K, int ldb, int b_kn,
                                      int out_dtype, int flag_base, unsigned epoch, unsigned flag_target, int num_sms,
                                      void* stream) {
    int rc = check_shapes(M, N, K, ldx, ldb, N, out_dtype);
    if (rc) return rc;
    CCMPI_REQUIRE(comms && nvr >= 1 && nvr <= CCMPI_MAX_VRANKS, "nvr=%d out of range", nvr);
    const int n = comms[0].size;
    CCMPI_REQUIRE(nvr == 1 || nvr == n, "emulation needs exactly one virtual rank per member (nvr=%d, size=%d)", nvr, n);
    CCMPI_REQUIRE(use_2cta(), "the gather-GEMM-gather kernel exists for the CTA-pair GEMM only (unset CCMPI_GEMM_1CTA)");
    const int m_blocks = (M + BM - 1) / BM;
    CCMPI_REQUIRE(flag_base >= 0 && flag_base + n * m_blocks <= CCMPI_PAD_USER_WORDS,
                  "fused op needs %d flag words at base %d, the pad has %d", n * m_blocks, flag_base, CCMPI_PAD_USER_WORDS);
    CCMPI_REQUIRE(K % n == 0 && (K / n) % BK == 0, "K/n = %d must be a multiple of %d", K / n, BK);
    CCMPI_REQUIRE((x_off | xfull_off | yfull_off) % 16 == 0, "symmetric offsets must be 16-byte aligned");
    const int chunk = K / n;
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const int budget = (sms / nvr) & ~1;
    CCMPI_REQUIRE(budget >= 2, "not enough SMs for %d emulated ranks", nvr);
    const int BN = pick_bn(M, N, budget);
    const int tiles = 2 * (((m_blocks + 1) / 2) * ((N + BN - 1) / BN));
    const int ctas = budget < tiles ? budget : tiles;
    CCMPI_REQUIRE(ctas <= CCMPI_MAX_BLOCKS, "grid of %d CTAs exceeds the %d barrier slots", ctas, CCMPI_MAX_BLOCKS);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        const ccmpi_dev_comm& c = comms[v];
        rc = fill_tmaps(rp, c.heap[c.rank] + xfull_off, c.heap[c.rank] + x_off, B[v], M, N, K, K, ldx, ldb, false,
                        b_kn != 0, BN / 2, M, chunk);
        if (rc) return rc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = c;
        rp.off0 = x_off; rp.off1 = xfull_off; rp.off2 = yfull_off;
        rp.chunk = chunk;
        rp.flag_base = flag_base;
        rp.epoch = flag_target ? flag_target : epoch * (unsigned)ctas;   // see ccmpi_fused_ag_gemm
        rp.aux0 = ldx;
        rp.aux1 = 0;         // bulk-copy push: the staging rows exist in this kernel
        rp.aux2 = 1;
    }
#define CCMPI_AGAG_CASE(bn, bmn)                                                                                  \
    if (BN == bn && (b_kn != 0) == bmn) {                                                                         \
        if (out_dtype == CCMPI_BF16)                                                                              \
            return launch2<bn, false, bmn, __nv_bfloat16, AgKAgOutPolicy>(lp, ctas * nvr, (cudaStream_t)stream, nvr > 1); \
        return launch2<bn, false, bmn, float, AgKAgOutPolicy>(lp, ctas * nvr, (cudaStream_t)stream, nvr > 1);     \
    }
    CCMPI_AGAG_CASE(256, false)
    CCMPI_AGAG_CASE(256, true)
    CCMPI_AGAG_CASE(128, false)
    CCMPI_AGAG_CASE(128, true)
#undef CCMPI_AGAG_CASE
    ccmpi::set_error("no CTA-pair GEMM instantiation for BN=%d", BN);
    return CCMPI_ERR_UNSUPPORTED;
}
