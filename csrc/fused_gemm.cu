// Host entry points of the fused GEMM + collective kernels (device side: fused_policies.cuh).
//
// Each entry takes `nvr` device communicators (1 in production; >1 = emulated ranks inside one
// cooperative launch on one GPU, used by the single-GPU tests) and, per rank, the local operand
// pointers. Buffers that peers touch are given as symmetric-heap offsets.
#include <cstdlib>

#include "fused_policies.cuh"
#include "gemm_host.h"

using namespace gemm;

namespace {

int common_checks(const ccmpi_dev_comm* comms, int nvr, int flag_words_needed, int flag_base) {
    CCMPI_REQUIRE(comms && nvr >= 1 && nvr <= CCMPI_MAX_VRANKS, "nvr=%d out of range", nvr);
    CCMPI_REQUIRE(nvr == 1 || nvr == comms[0].size, "emulation needs exactly one virtual rank per member (nvr=%d, size=%d)",
                  nvr, comms[0].size);
    CCMPI_REQUIRE(flag_base >= 0 && flag_base + flag_words_needed <= CCMPI_PAD_USER_WORDS,
                  "fused op needs %d flag words at base %d, the pad has %d: use smaller M or a bigger tile", flag_words_needed,
                  flag_base, CCMPI_PAD_USER_WORDS);
    return CCMPI_OK;
}

int grid_for(int nvr, int sms, int tiles, int extra) {
    int per_rank = sms / nvr;
    if (per_rank > tiles + extra) per_rank = tiles + extra;
    return per_rank;
}

}  // namespace

// all-gather -> GEMM (K-gather): C_r[M][N] = [x_0 | ... | x_{n-1}] * B_r, x_r = rank r's [M][K/n] shard
extern "C" int ccmpi_fused_ag_gemm(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                                   const void* const* B, void* const* C, int M, int N, int K, int ldb, int ldc,
                                   int b_kn, int out_dtype, int flag_base, unsigned epoch, unsigned flag_target,
                                   int num_comm_ctas, int num_sms, int use_mc, void* stream) {
    return ccmpi_fused_ag_gemm_ex(comms, nvr, x_off, ldx, xfull_off, B, C, M, N, K, ldb, ldc, b_kn, out_dtype, flag_base,
                                  epoch, flag_target, 0, nullptr, nullptr, 0, nullptr, num_comm_ctas, num_sms, use_mc, stream);
}

// ... with a fused elementwise epilogue (EpiOp; epi_in / epi_out2: one pointer per rank, or NULL) and a device trace
extern "C" int ccmpi_fused_ag_gemm_ex(const ccmpi_dev_comm* comms, int nvr, size_t x_off, int ldx, size_t xfull_off,
                                      const void* const* B, void* const* C, int M, int N, int K, int ldb, int ldc,
                                      int b_kn, int out_dtype, int flag_base, unsigned epoch, unsigned flag_target,
                                      int epi_op, const void* const* epi_in, void* const* epi_out2, long long epi_ld,
                                      void* trace, int num_comm_ctas, int num_sms, int use_mc, void* stream) {
    int rc = check_shapes(M, N, K, ldx, ldb, ldc, out_dtype);
    if (rc) return rc;
    const int n = comms ? comms[0].size : 0;
    const int m_blocks = (M + BM - 1) / BM;
    rc = common_checks(comms, nvr, n * m_blocks, flag_base);
    if (rc) return rc;
    CCMPI_REQUIRE(K % n == 0 && (K / n) % BK == 0, "K/n = %d must be a multiple of %d", K / n, BK);
    CCMPI_REQUIRE((x_off | xfull_off) % 16 == 0, "symmetric offsets must be 16-byte aligned");
    const int chunk = K / n;
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const bool pairs = use_2cta();
    const int ncomm = 0;  // the gather is pushed by every CTA's epilogue warps (AgKPolicy::prologue)
    int per_rank_budget = sms / nvr;
    if (pairs) per_rank_budget &= ~1;
    CCMPI_REQUIRE(per_rank_budget > ncomm, "not enough SMs (%d per rank) for %d communication CTAs", per_rank_budget, ncomm);
    const int BN = pick_bn(M, N, per_rank_budget - ncomm);
    const int tiles = pairs ? 2 * (((m_blocks + 1) / 2) * ((N + BN - 1) / BN)) : m_blocks * ((N + BN - 1) / BN);
    int gemm_ctas = per_rank_budget - ncomm;
    if (gemm_ctas > tiles) gemm_ctas = tiles;
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        const ccmpi_dev_comm& c = comms[v];
        const void* xfull = c.heap[c.rank] + xfull_off;
        const void* xloc = c.heap[c.rank] + x_off;
        rc = fill_tmaps(rp, xfull, xloc, B[v], M, N, K, K, ldx, ldb, false, b_kn != 0, pairs ? BN / 2 : BN, M, chunk);
        if (rc) return rc;
        rp.C = C[v];
        rp.ldc = ldc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = c;
        rp.off0 = x_off; rp.off1 = xfull_off;
        rp.chunk = chunk;
        rp.flag_base = flag_base;
        // A ready flag counts the pushing CTAs of its source rank. flag_target (what the flags must
        // reach for THIS launch: the caller's running sum of CTA counts, ccmpi_fused_ag_ctas) lets the
        // CTA count differ from launch to launch; 0 = every launch on these flags has this CTA count.
        rp.epoch = flag_target ? flag_target : epoch * (unsigned)gemm_ctas;
        rp.num_comm_ctas = ncomm;
        rp.aux0 = ldx;
        static const int env_push_first = [] { const char* e = getenv("CCMPI_AG_PUSH_FIRST"); return e ? atoi(e) : 0; }();
        rp.aux1 = (use_mc ? 1 : 0) | (env_push_first ? 4 : 0);
        rp.aux2 = 1;
        rc = set_epilogue(rp, epi_op, epi_in ? epi_in[v] : nullptr, epi_out2 ? epi_out2[v] : nullptr, epi_ld, N, out_dtype, pairs);
        if (rc) return rc;
        rp.trace = pairs ? (unsigned long long*)trace : nullptr;
    }
    if (pairs)
        return dispatch2<AgKPolicy>(lp, BN, false, b_kn != 0, out_dtype, (gemm_ctas + ncomm) * nvr, (cudaStream_t)stream, nvr > 1);
    return dispatch<AgKPolicy>(lp, BN, false, b_kn != 0, out_dtype, (gemm_ctas + ncomm) * nvr, (cudaStream_t)stream, nvr > 1);
}

// CTAs per rank that a launch of an all-gather -> GEMM kernel will use (= contributions each ready flag
// receives from one source rank): kind 0 K-gather (ccmpi_fused_ag_gemm, ccmpi_fused_ag_gemm_ag), 1 M-gather.
extern "C" int ccmpi_fused_ag_ctas(int kind, int M, int N, int nvr, int num_sms) {
    CCMPI_REQUIRE(nvr >= 1 && M > 0 && N > 0, "bad arguments");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    int budget = sms / nvr;
    const bool pairs = use_2cta() || kind == 1 || kind == 2;
    if (pairs) budget &= ~1;
    const int BN = kind == 2 ? 128 : pick_bn(M, N, budget);     // kind 2: block-scaled K-gather (256 x 128 tiles)
    const int m_blocks = (M + BM - 1) / BM;
    const int nb = (N + BN - 1) / BN;
    const int tiles = kind == 1 ? 2 * ((m_blocks / 2) * nb) : (pairs ? 2 * (((m_blocks + 1) / 2) * nb) : m_blocks * nb);
    CCMPI_REQUIRE(kind >= 0 && kind <= 2, "unknown kernel kind %d", kind);
    return budget < tiles ? budget : tiles;
}

namespace {
// shared by reduce-scatter and all-reduce flavours
template <typename Policy>
int gemm_scatter(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B, size_t slots_off,
                 void* const* C, size_t out_off, int M, int N, int K, int lda, int ldb, int ldc, int a_km, int b_kn,
                 int out_dtype, int flag_base, unsigned epoch, int num_sms, int use_mc, void* stream,
                 const void* const* residual = nullptr, long long res_ld = 0, void* trace = nullptr) {
    int rc = check_shapes(M, N, K, lda, ldb, ldc, out_dtype);
    if (rc) return rc;
    const int n = comms ? comms[0].size : 0;
    CCMPI_REQUIRE(n >= 1 && N % n == 0, "N = %d must be divisible by the communicator size %d", N, n);
    const int chunk = N / n;
    CCMPI_REQUIRE(chunk % 128 == 0, "N/n = %d must be a multiple of 128", chunk);
    const int BN = (chunk % 256 == 0) ? 256 : 128;
    const int m_blocks = (M + BM - 1) / BM;
    rc = common_checks(comms, nvr, m_blocks * (chunk / BN) * n, flag_base);
    if (rc) return rc;
    CCMPI_REQUIRE((slots_off | out_off) % 16 == 0, "symmetric offsets must be 16-byte aligned");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const bool pairs = use_2cta();
    const int tiles = pairs ? 2 * (((m_blocks + 1) / 2) * (N / BN)) : m_blocks * (N / BN);
    int per_rank = grid_for(nvr, sms, tiles, 0);
    if (pairs) per_rank &= ~1;
    CCMPI_REQUIRE(per_rank >= (pairs ? 2 : 1), "not enough SMs for %d emulated ranks", nvr);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        rc = fill_tmaps(rp, A[v], A[v], B[v], M, N, K, lda, lda, ldb, a_km != 0, b_kn != 0, pairs ? BN / 2 : BN,
                        a_km ? K : M, a_km ? M : K);
        if (rc) return rc;
        rp.C = C ? C[v] : nullptr;
        rp.ldc = ldc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = comms[v];
        rp.off1 = slots_off; rp.off2 = out_off;
        rp.chunk = chunk;
        rp.flag_base = flag_base;
        rp.epoch = epoch;
        rp.aux0 = BN;
        rp.aux1 = use_mc;
        if (residual) {     // added to the reduced tile where it is finished (RsNPolicy::own_store / RsNSlotPolicy::reduce_tile)
            CCMPI_REQUIRE(out_dtype == CCMPI_BF16 && residual[v] != nullptr && res_ld % 8 == 0 && res_ld >= chunk,
                          "reduce-scatter residual: bf16 output, non-null pointer, row stride %% 8 == 0 and >= N/n");
            rp.epi_op = EPI_ADD;
            rp.epi_in = residual[v];
            rp.epi_ld = res_ld;
        }
        rp.trace = pairs ? (unsigned long long*)trace : nullptr;
    }
    if (pairs) return dispatch2<Policy>(lp, BN, a_km != 0, b_kn != 0, out_dtype, per_rank * nvr, (cudaStream_t)stream, nvr > 1);
    return dispatch<typename Policy::SingleCta>(lp, BN, a_km != 0, b_kn != 0, out_dtype, per_rank * nvr, (cudaStream_t)stream, nvr > 1);
}
}  // namespace

// GEMM -> reduce-scatter (N-scatter): C_r[M][N/n] = sum over ranks s of (A_s * B_s)[:, r-th column block]
extern "C" int ccmpi_fused_gemm_rs(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                                   size_t slots_off, void* const* C, int M, int N, int K, int lda, int ldb, int ldc,
                                   int a_km, int b_kn, int out_dtype, int flag_base, unsigned epoch, int num_sms,
                                   void* stream) {
    CCMPI_REQUIRE(C != nullptr, "gemm_rs needs an output pointer per rank");
    return gemm_scatter<RsNPolicy>(comms, nvr, A, B, slots_off, C, 0, M, N, K, lda, ldb, ldc, a_km, b_kn, out_dtype,
                                   flag_base, epoch, num_sms, 0, stream);
}

// ... plus `residual` ([M][N/n] bf16 per rank, row stride res_ld) added to the reduced result, and a device trace
extern "C" int ccmpi_fused_gemm_rs_ex(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                                      size_t slots_off, void* const* C, int M, int N, int K, int lda, int ldb, int ldc,
                                      int a_km, int b_kn, int out_dtype, int flag_base, unsigned epoch,
                                      const void* const* residual, long long res_ld, void* trace, int num_sms, void* stream) {
    CCMPI_REQUIRE(C != nullptr, "gemm_rs needs an output pointer per rank");
    return gemm_scatter<RsNPolicy>(comms, nvr, A, B, slots_off, C, 0, M, N, K, lda, ldb, ldc, a_km, b_kn, out_dtype,
                                   flag_base, epoch, num_sms, 0, stream, residual, res_ld, trace);
}

// GEMM -> all-reduce: out[M][N] (symmetric, row stride N) = sum over ranks s of A_s * B_s, on every rank
extern "C" int ccmpi_fused_gemm_ar(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                                   size_t slots_off, size_t out_off, int M, int N, int K, int lda, int ldb, int a_km,
                                   int b_kn, int out_dtype, int flag_base, unsigned epoch, int num_sms, int use_mc,
                                   void* stream) {
    return gemm_scatter<ArPolicy>(comms, nvr, A, B, slots_off, nullptr, out_off, M, N, K, lda, ldb, N, a_km, b_kn,
                                  out_dtype, flag_base, epoch, num_sms, use_mc, stream);
}

// GEMM -> all-gather of the output: y_full[M][N*n] (symmetric) gets this rank's A*B at column block `rank`
extern "C" int ccmpi_fused_gemm_ag(const ccmpi_dev_comm* comms, int nvr, const void* const* A, const void* const* B,
                                   size_t yfull_off, int M, int N, int K, int lda, int ldb, int a_km, int b_kn,
                                   int out_dtype, int num_sms, int use_mc, void* stream) {
    int rc = check_shapes(M, N, K, lda, ldb, N, out_dtype);
    if (rc) return rc;
    rc = common_checks(comms, nvr, 0, 0);
    if (rc) return rc;
    CCMPI_REQUIRE(yfull_off % 16 == 0, "symmetric offsets must be 16-byte aligned");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count());
    const bool pairs = use_2cta();
    const int BN = pick_bn(M, N, sms / nvr);
    const int m_blocks = (M + BM - 1) / BM;
    const int tiles = pairs ? 2 * (((m_blocks + 1) / 2) * ((N + BN - 1) / BN)) : m_blocks * ((N + BN - 1) / BN);
    int per_rank = grid_for(nvr, sms, tiles, 0);
    if (pairs) per_rank &= ~1;
    CCMPI_REQUIRE(per_rank <= CCMPI_MAX_BLOCKS, "grid of %d CTAs exceeds the %d barrier slots", per_rank, CCMPI_MAX_BLOCKS);
    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int v = 0; v < nvr; ++v) {
        RankParams& rp = lp.r[v];
        rc = fill_tmaps(rp, A[v], A[v], B[v], M, N, K, lda, lda, ldb, a_km != 0, b_kn != 0, pairs ? BN / 2 : BN,
                        a_km ? K : M, a_km ? M : K);
        if (rc) return rc;
        rp.M = M; rp.N = N; rp.K = K;
        rp.comm = comms[v];
        rp.off1 = yfull_off;
        rp.aux1 = use_mc;
    }
    if (pairs) return dispatch2<AgOutPolicy>(lp, BN, a_km != 0, b_kn != 0, out_dtype, per_rank * nvr, (cudaStream_t)stream, nvr > 1);
    return dispatch<AgOutPolicy>(lp, BN, a_km != 0, b_kn != 0, out_dtype, per_rank * nvr, (cudaStream_t)stream, nvr > 1);
}
