// The persistent tcgen05 GEMM kernel body, parameterised by a Policy (see gemm_tcgen05.cuh).
#pragma once
#include "act.cuh"
#include "gemm_tcgen05.cuh"

namespace gemm {

// ------------------------------------------------------------------------------------------------
// Output conversion helpers: 32 fp32 accumulator values of one row -> global memory
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16(uint32_t lo_bits, uint32_t hi_bits) {
    __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(lo_bits), __uint_as_float(hi_bits));
    return *reinterpret_cast<uint32_t*>(&h);
}

// Store up to 32 columns of one row. `ncols` (multiple of 8 for bf16, 4 for fp32) handles the
// right edge. dst points at (row, col0).
template <typename OutT>
__device__ __forceinline__ void store_row32(OutT* dst, const uint32_t (&v)[32], int ncols, bool accumulate) {
    if constexpr (sizeof(OutT) == 2) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (j * 8 < ncols) {
                uint4 q;
                q.x = pack_bf16(v[8 * j + 0], v[8 * j + 1]);
                q.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
                q.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]);
                q.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
                ptx::st_v4(dst + 8 * j, q);
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (j * 4 < ncols) {
                uint4 q = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                if (accumulate) {
                    const uint4 o = ptx::ld_v4(dst + 4 * j);
                    q.x = __float_as_uint(__uint_as_float(q.x) + __uint_as_float(o.x));
                    q.y = __float_as_uint(__uint_as_float(q.y) + __uint_as_float(o.y));
                    q.z = __float_as_uint(__uint_as_float(q.z) + __uint_as_float(o.z));
                    q.w = __float_as_uint(__uint_as_float(q.w) + __uint_as_float(o.w));
                }
                ptx::st_v4(dst + 4 * j, q);
            }
        }
    }
}

// 128 epilogue threads only (named barrier 1); never stalls the producer / MMA warps.
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------
// Plain policy: C = A*B, row-major C, grouped raster, no communication.
// ------------------------------------------------------------------------------------------------
struct PlainPolicy {
    static constexpr bool kCommCtas = false;
    static constexpr bool kStagedStore = false;   // CTA-pair kernel: stage tiles in smem + bulk copies
    static constexpr bool kNeedsStaging = false;  // CTA-pair kernel: reserve the per-lane staging rows
    static constexpr bool kOwnTileDirect = false; // CTA-pair kernel, staged policies: some tiles are finished in place (RsNPolicy)

    // Background work: something the epilogue warps do while no accumulator is ready (pushing a
    // gathered operand, reducing scattered tiles). `State` carries its progress.
    struct State {};
    static constexpr bool kBackground = false;
    // `stg` = this lane's 256-byte staging row (0 if the kernel has none), `wbar` = an mbarrier
    // private to this epilogue warp (nullptr if none)
    __device__ static void init_state(State& st, const RankParams& p, int cta, int nctas, int epi_tid, uint32_t stg, uint64_t* wbar) {}
    // one bounded slice of background work; false = nothing (more) to do right now.
    // `bcast` is a shared-memory word the 128 epilogue threads may use to agree on a decision.
    template <typename OutT>
    __device__ static bool background(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) { return false; }
    static constexpr bool kPushFirst = false;

    __device__ static int num_tiles(const RankParams& p, int mb, int nb) { return mb * nb; }
    __device__ static TileCoord tile(const RankParams& p, int t, int mb, int nb) { return raster(t, mb, nb); }
    // producer: K-block `it` of tile `tc` -> coordinates + operand source
    __device__ static void kblock(const RankParams& p, const TileCoord& tc, int it, int& k0, const CUtensorMap*& ta,
                                  int& a_k0, int& a_m0) {
        k0 = it * BK;
        ta = &p.tmap_a;
        a_k0 = k0;
        a_m0 = tc.m_blk * BM;
    }
    // CTA-pair kernel: K is walked in `num_kchunks` equal chunks and kchunk() is called ONCE per (tile, chunk) --
    // it may wait for the chunk to arrive (`wait`; false for the zero-filled phantom m-block of an odd tail) and
    // says where the chunk's A rows come from; the per-K-block loop then only adds BK. (The producer thread has
    // ~270 ns per K-block at full MMA rate: the divisions a per-K-block kblock() needs to find its chunk were
    // measured at ~150 ns per call with the device trace, i.e. they starved the tensor cores.)
    __device__ static int num_kchunks(const RankParams& p) { return 1; }
    __device__ static void kchunk(const RankParams& p, const TileCoord& tc, int ci, bool wait, int& k_base,
                                  const CUtensorMap*& ta, int& a_k_base, int& a_m0) {
        k_base = 0;
        ta = &p.tmap_a;
        a_k_base = 0;
        a_m0 = tc.m_blk * BM;
    }
    // K coordinate of iteration `it` (what kblock would use for B), without any waiting
    __device__ static int kcoord(const RankParams& p, int it) { return it * BK; }
    template <typename OutT>
    __device__ static void store(const RankParams& p, const TileCoord& tc, int row, int col0, const uint32_t (&v)[32]) {
        if (row < p.M) {
            const int ncols = min(32, p.N - col0);
            if (ncols > 0)
                store_row32<OutT>(reinterpret_cast<OutT*>(p.C) + (long long)row * p.ldc + col0, v, ncols,
                                  p.accumulate != 0);
        }
    }
    // ---- fused elementwise epilogues (RankParams::epi_op; bf16 C, N % 32 == 0 checked on the host) ----
    // Policies that keep PlainPolicy's store (plain GEMM, all-gather -> GEMM) inherit these; policies with
    // their own store set kEpiOps = false.
    static constexpr bool kEpiOps = true;
    __device__ static bool epi_has_input(const RankParams& p) { return p.epi_op == EPI_MUL_DGELU || p.epi_op == EPI_ADD; }
    // this lane's 32 bf16 of the second operand at (row, col0): issued one chunk ahead of its use
    __device__ static void epi_load(const RankParams& p, int row, int col0, uint4 (&e)[4]) {
        if (row < p.M && col0 < p.N) {
            const __nv_bfloat16* s = reinterpret_cast<const __nv_bfloat16*>(p.epi_in) + (long long)row * p.epi_ld + col0;
#pragma unroll
            for (int j = 0; j < 4; ++j) e[j] = ptx::ld_v4(s + 8 * j);
        }
    }
    template <typename OutT>
    __device__ static void store_epi(const RankParams& p, const TileCoord& tc, int row, int col0, uint32_t (&v)[32],
                                     const uint4 (&e)[4]) {
        if (row >= p.M || col0 >= p.N) return;
        OutT* dst = reinterpret_cast<OutT*>(p.C) + (long long)row * p.ldc + col0;
        if (p.epi_op == EPI_GELU_DUAL) {
            store_row32<OutT>(dst, v, 32, false);
            // the activation is computed from the ROUNDED pre-activation, i.e. from what backward will read
#pragma unroll
            for (int i = 0; i < 32; ++i)
                v[i] = __float_as_uint(act::gelu(__bfloat162float(__float2bfloat16_rn(__uint_as_float(v[i])))));
            store_row32<OutT>(reinterpret_cast<OutT*>(p.epi_out2) + (long long)row * p.epi_ld + col0, v, 32, false);
            return;
        }
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&e[0]);
        if (p.epi_op == EPI_MUL_DGELU) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 u = __bfloat1622float2(h[i]);
                v[2 * i] = __float_as_uint(__uint_as_float(v[2 * i]) * act::gelu_grad(u.x));
                v[2 * i + 1] = __float_as_uint(__uint_as_float(v[2 * i + 1]) * act::gelu_grad(u.y));
            }
        } else {   // EPI_ADD
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 r = __bfloat1622float2(h[i]);
                v[2 * i] = __float_as_uint(__uint_as_float(v[2 * i]) + r.x);
                v[2 * i + 1] = __float_as_uint(__uint_as_float(v[2 * i + 1]) + r.y);
            }
        }
        store_row32<OutT>(dst, v, 32, false);
    }
    // called by all 128 epilogue threads before their first tile
    __device__ static void prologue(const RankParams& p, int cta, int nctas, int epi_tid) {}
    // called by all 128 epilogue threads after every tile
    __device__ static void tile_done(const RankParams& p, const TileCoord& tc, int epi_tid) {}
    // called by all epilogue threads once the CTA has no more tiles
    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {}
    // body of a communication CTA (only if kCommCtas)
    __device__ static void comm_cta(const RankParams& p, int comm_cta, int ncomm) {}
};

// ------------------------------------------------------------------------------------------------
// The kernel
// ------------------------------------------------------------------------------------------------
template <int BN, bool A_MN, bool B_MN, typename OutT, typename Policy>
__global__ void __launch_bounds__(NUM_THREADS, 1) gemm_kernel(const __grid_constant__ Launch lp) {
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    Smem<BN> sm(smem_raw);

    // virtual-rank decomposition (nvr == 1 in production)
    const int nvr = lp.nvr;
    const int vr = blockIdx.x % nvr;
    int cta = blockIdx.x / nvr;
    const int nctas_all = gridDim.x / nvr;
    const RankParams& p = lp.r[vr];

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if constexpr (Policy::kCommCtas) {
        // The LEADING CTAs of each rank move data (lowest block indices are scheduled first, so
        // they are resident even if something else holds a few SMs); they never touch TMEM.
        const int ncomm = p.num_comm_ctas;
        if (cta < ncomm) {
            Policy::comm_cta(p, cta, ncomm);
            return;
        }
        cta -= ncomm;
    }
    const int nctas = nctas_all - (Policy::kCommCtas ? p.num_comm_ctas : 0);

    const int m_blocks = (p.M + BM - 1) / BM;
    const int n_blocks = (p.N + BN - 1) / BN;
    const int k_blocks = (p.K + BK - 1) / BK;
    const int ntiles = Policy::num_tiles(p, m_blocks, n_blocks);

    // ---- one-time setup ----
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&p.tmap_a);
        ptx::prefetch_tmap(&p.tmap_a2);
        ptx::prefetch_tmap(&p.tmap_b);
#pragma unroll
        for (int s = 0; s < C::STAGES; ++s) {
            ptx::mbar_init(&sm.full[s], 1);
            ptx::mbar_init(&sm.empty[s], 1);
        }
        ptx::mbar_init(&sm.tmem_full[0], 1);
        ptx::mbar_init(&sm.tmem_full[1], 1);
        ptx::mbar_init(&sm.tmem_empty[0], 4);  // one arrival per epilogue warp
        ptx::mbar_init(&sm.tmem_empty[1], 4);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc<C::TMEM_COLS>(sm.tmem_base);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *sm.tmem_base;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = cta; t < ntiles; t += nctas) {
                const TileCoord tc = Policy::tile(p, t, m_blocks, n_blocks);
                for (int it = 0; it < k_blocks; ++it) {
                    int k0, a_k0, a_m0;
                    const CUtensorMap* ta;
                    Policy::kblock(p, tc, it, k0, ta, a_k0, a_m0);  // may wait on gather flags
                    ptx::mbar_wait(&sm.empty[stage], phase ^ 1, 100 + stage);
                    ptx::mbar_arrive_expect_tx(&sm.full[stage], C::STAGE_BYTES);
                    load_operand<A_MN, BM>(sm.a(stage), ta, &sm.full[stage], a_k0, a_m0);
                    load_operand<B_MN, BN>(sm.b(stage), &p.tmap_b, &sm.full[stage], k0, tc.n_blk * BN);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::umma_idesc(1, 1, A_MN, B_MN, BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int local = 0;
            for (int t = cta; t < ntiles; t += nctas, ++local) {
                const int acc = local & 1;
                const uint32_t acc_phase = (local >> 1) & 1;
                ptx::mbar_wait(&sm.tmem_empty[acc], acc_phase ^ 1, 200 + acc);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int it = 0; it < k_blocks; ++it) {
                    ptx::mbar_wait(&sm.full[stage], phase, 300 + stage);
                    ptx::tc_fence_after();
                    const uint32_t a_addr = ptx::smem_u32(sm.a(stage));
                    const uint32_t b_addr = ptx::smem_u32(sm.b(stage));
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
                        ptx::umma_f16_ss(tmem_d, operand_desc<A_MN>(a_addr, ks), operand_desc<B_MN>(b_addr, ks), idesc,
                                         (it | ks) != 0 ? 1u : 0u);
                    }
                    ptx::umma_commit(&sm.empty[stage]);  // smem slot reusable once these MMAs retire
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit(&sm.tmem_full[acc]);    // accumulator complete -> epilogue
            }
        }
        __syncwarp();
    } else {
        // ===================== epilogue warps =====================
        const int q = warp & 3;                  // TMEM lane quarter this warp may access
        const int epi_tid = (warp - EPI_WARP0) * 32 + lane;
        int local = 0;
        typename Policy::State st;
        volatile uint32_t* bcast = sm.tmem_base + 1;
        Policy::init_state(st, p, cta, nctas, epi_tid, 0u, nullptr);
        Policy::prologue(p, cta, nctas, epi_tid);
        if constexpr (Policy::kBackground) {
            // single-CTA kernel: no interleaving, pushes are finished up front
            if (Policy::kPushFirst)
                while (Policy::template background<OutT>(p, st, cta, nctas, epi_tid, bcast)) {}
        }
        for (int t = cta; t < ntiles; t += nctas, ++local) {
            const TileCoord tc = Policy::tile(p, t, m_blocks, n_blocks);
            const int acc = local & 1;
            const uint32_t acc_phase = (local >> 1) & 1;
            ptx::mbar_wait(&sm.tmem_full[acc], acc_phase, 400 + acc);
            ptx::tc_fence_after();
            const int row = tc.m_blk * BM + q * 32 + lane;
            const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t v[32];
                ptx::tmem_ld_32x32b_x32(taddr + c, v);
                ptx::tmem_ld_wait();
                if (c + 32 >= BN) {
                    // all TMEM reads of this accumulator stage are done: hand it back early
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(&sm.tmem_empty[acc]);
                }
                Policy::template store<OutT>(p, tc, row, tc.n_blk * BN + c, v);
            }
            Policy::tile_done(p, tc, epi_tid);
        }
        Policy::template tail<OutT>(p, st, cta, nctas, epi_tid, bcast);
    }

    // ---- teardown ----
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<C::TMEM_COLS>(tmem_base);
    }
}

}  // namespace gemm
