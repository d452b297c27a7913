// Token-axis (GEMM M) variants of the fused kernels -- the Megatron "sequence parallel" form:
//
//   AgMPolicy   all-gather -> GEMM with the gather on M:  x_r is rank r's [M/n][K] block of rows
//   RsMPolicy   GEMM -> reduce-scatter with the scatter on M: rank r keeps rows [r*M/n, (r+1)*M/n)
//
// The reference only ever shards the FEATURE axis (func_impl.py:90,108,182), which is what
// fused_policies.cuh implements and what all default paths use. These two policies reuse the same
// machinery (background push on the bulk-copy engine, staged bulk epilogue, per-tile flags,
// opportunistic rank-ordered reduction); only the index arithmetic differs.
//
// Run on B200s in round 2: emulated ranks on one GPU and real ranks on 2 / 4 / 8 GPUs
// (tests/test_gpu_seq.py, profiles/r2_seq_bench_2gpu.jsonl). CTA-pair kernel only.
#pragma once
#include "fused_policies.cuh"

namespace gemm {

// ================================================================================================
//   off0 : symmetric offset of x_local [Mr][K]   (Mr = M/n rows, row stride aux0 elements)
//   off1 : symmetric offset of x_full  [M][K]
//   chunk: Mr (rows per rank; a multiple of 256 so that CTA pairs never straddle two owners)
//   aux2 : CTAs per rank (contributions per ready flag and launch)
//   tmap_a -> x_full, tmap_a2 -> x_local
// ================================================================================================
struct AgMPolicy : PlainPolicy {
    static constexpr bool kBackground = true;
    static constexpr bool kPushFirst = true;
    static constexpr bool kNeedsStaging = true;

    __device__ static int mbpr(const RankParams& p) { return p.chunk / BM; }   // m-blocks per rank
    __device__ static int ready_idx(const RankParams& p, int src, int lm) { return p.flag_base + src * mbpr(p) + lm; }

    // Owner-major order: ALL tiles of this rank's own rows first (they need no remote data), then the rows of
    // rank+1, rank+2, ...; the grouped raster runs inside one owner's block of rows. (Rotating one global raster
    // is not enough: a raster group of 8 m-pairs spans several owners as soon as a rank has fewer than 8 pairs.)
    __device__ static TileCoord tile(const RankParams& p, int t, int mb, int nb) {
        const int n = p.comm.size;
        const int per_owner = (mb / n) * nb;
        const int o = t / per_owner;
        TileCoord tc = raster(t - o * per_owner, mb / n, nb);
        tc.m_blk += ((p.comm.rank + o) % n) * (mb / n);
        return tc;
    }

    __device__ static void kblock(const RankParams& p, const TileCoord& tc, int it, int& k0, const CUtensorMap*& ta,
                                  int& a_k0, int& a_m0) {
        const ccmpi_dev_comm& c = p.comm;
        k0 = it * BK;
        a_k0 = k0;
        const int src = tc.m_blk / mbpr(p);
        const int lm = tc.m_blk - src * mbpr(p);
        if (src == c.rank) {
            ta = &p.tmap_a2;
            a_m0 = lm * BM;
        } else {
            if (it == 0) {
                ptx::wait_flag_ge_sys(user_flag(c, c.rank, ready_idx(p, src, lm)), p.epoch * (unsigned)p.aux2, 520 + src);
                ptx::fence_proxy_async_all();
            }
            ta = &p.tmap_a;
            a_m0 = tc.m_blk * BM;
        }
    }

    __device__ static void kchunk(const RankParams& p, const TileCoord& tc, int ci, bool wait, int& k_base,
                                  const CUtensorMap*& ta, int& a_k_base, int& a_m0) {
        const ccmpi_dev_comm& c = p.comm;
        k_base = 0;
        a_k_base = 0;
        const int src = tc.m_blk / mbpr(p);
        const int lm = tc.m_blk - src * mbpr(p);
        if (src == c.rank) {
            ta = &p.tmap_a2;
            a_m0 = lm * BM;
        } else {
            if (wait) {
                ptx::wait_flag_ge_sys(user_flag(c, c.rank, ready_idx(p, src, lm)), p.epoch * (unsigned)p.aux2, 520 + src);
                ptx::fence_proxy_async_all();
            }
            ta = &p.tmap_a;
            a_m0 = tc.m_blk * BM;
        }
    }

    struct State {
        long long base;
        int flagged, m_blocks;
        uint32_t stg;
        uint64_t* wbar;
        uint32_t phase;
    };
    __device__ static void init_state(State& st, const RankParams& p, int cta, int nctas, int epi_tid, uint32_t stg, uint64_t* wbar) {
        st.base = (long long)cta * EPI_THREADS;
        st.flagged = 0;
        st.m_blocks = p.comm.size == 1 ? 0 : mbpr(p);
        st.stg = stg;
        st.wbar = wbar;
        st.phase = 0;
        if ((epi_tid & 31) == 0) {
            ptx::mbar_init(wbar, 32);
            ptx::fence_mbar_init();
        }
        __syncwarp();
    }

    // one 256-byte piece per epilogue lane per slice: local rows -> every member's x_full rows
    template <typename OutT>
    __device__ static bool background(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        if (st.flagged >= st.m_blocks) return false;
        const ccmpi_dev_comm& c = p.comm;
        const int row_bytes = p.K * 2;
        const int PB = (row_bytes % 256 == 0) ? 256 : 128;
        const int ppr = row_bytes / PB;
        const long long total = (long long)p.chunk * ppr;
        const long long per_blk = (long long)BM * ppr;
        const long long stride = (long long)nctas * EPI_THREADS;
        const uint8_t* x = c.heap[c.rank] + p.off0;
        const size_t src_ld = (size_t)p.aux0 * 2, dst_ld = (size_t)p.K * 2;
        const long long i = st.base + epi_tid;
        const bool ok = i < total;
        ptx::tma_store_wait_read<0>();
        ptx::mbar_arrive_expect_tx(st.wbar, ok ? (uint32_t)PB : 0u);
        size_t d = 0;
        if (ok) {
            const long long row = i / ppr;
            const int w = (int)(i - row * ppr);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(st.stg),
                         "l"(x + (size_t)row * src_ld + (size_t)w * PB), "r"(PB), "r"(ptx::smem_u32(st.wbar))
                         : "memory");
            d = p.off1 + ((size_t)c.rank * p.chunk + (size_t)row) * dst_ld + (size_t)w * PB;
        }
        ptx::mbar_wait(st.wbar, st.phase, 810);
        st.phase ^= 1;
        if (ok) {
#pragma unroll
            for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
                if (s < c.size) ptx::bulk_s2g(c.heap[(c.rank + s) % c.size] + d, st.stg, PB);
        }
        ptx::tma_store_commit();
        const long long next = st.base + stride;
        int done = (int)min((long long)st.m_blocks, next / per_blk);
        // flags go out per group of m-blocks (each costs a system-scope fence); a rank's share is only M/n rows
        // here, so small shares use a finer grain than the K-gather's 16
        const int granule = st.m_blocks >= 32 ? 16 : 4;
        done = (done / granule) * granule;
        if (next >= total) done = st.m_blocks;
        st.base = next;
        if (done <= st.flagged) return true;
        ptx::tma_store_wait<0>();
        ptx::fence_proxy_async_all();
        epi_bar_sync();
        if (epi_tid == 0) {
            ptx::fence_acq_rel_sys();
            for (int j = st.flagged; j < done; ++j)
                for (int s = 1; s < c.size; ++s)
                    ptx::red_relaxed_sys_add(user_flag(c, (c.rank + s) % c.size, ready_idx(p, c.rank, j)), 1u);
        }
        st.flagged = done;
        return true;
    }

    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        while (background<OutT>(p, st, cta, nctas, epi_tid, bcast)) {}
    }
};

// ================================================================================================
//   off1 : symmetric offset of the slot buffer [n sources][Mr][N] (OutT)
//   C    : this rank's reduced rows [Mr][N], row stride ldc
//   chunk: Mr;  aux0: BN
// ================================================================================================
struct RsMPolicy : PlainPolicy {
    static constexpr bool kEpiOps = false;   // own store: no fused elementwise epilogue
    static constexpr bool kStagedStore = true;
    static constexpr bool kNeedsStaging = true;
    static constexpr bool kBackground = true;

    __device__ static int mbpr(const RankParams& p) { return p.chunk / BM; }
    __device__ static int nblocks(const RankParams& p) { return p.N / p.aux0; }
    __device__ static int flag_idx(const RankParams& p, int lm, int n_blk, int src) {
        return p.flag_base + ((lm * nblocks(p) + n_blk) * p.comm.size + src);
    }

    // owners interleaved tile by tile (rank+1, ..., rank), like the N-scatter policy
    __device__ static TileCoord tile(const RankParams& p, int t, int mb, int nb) {
        const int n = p.comm.size;
        const int g = t % n, in_g = t / n;
        const int owner = (p.comm.rank + 1 + g) % n;
        TileCoord tc = raster(in_g, mb / n, nb);
        tc.m_blk += owner * (mb / n);
        return tc;
    }

    template <typename OutT>
    __device__ static void bulk_row(const RankParams& p, const TileCoord& tc, int row, int col0, uint32_t smem_row, int bytes) {
        const ccmpi_dev_comm& c = p.comm;
        const int owner = row / p.chunk;
        const int lrow = row - owner * p.chunk;
        OutT* slot = reinterpret_cast<OutT*>(c.heap[owner] + p.off1) + ((size_t)c.rank * p.chunk + lrow) * p.N + col0;
        ptx::bulk_s2g(slot, smem_row, bytes);
    }

    __device__ static void tile_done(const RankParams& p, const TileCoord& tc, int epi_tid) {
        epi_bar_sync();
        if (epi_tid == 0) {
            const ccmpi_dev_comm& c = p.comm;
            const int owner = tc.m_blk / mbpr(p);
            ptx::fence_acq_rel_sys();
            ptx::st_release_sys(user_flag(c, owner, flag_idx(p, tc.m_blk - owner * mbpr(p), tc.n_blk, c.rank)), p.epoch);
        }
    }

    struct State { int next; };
    __device__ static void init_state(State& st, const RankParams& p, int cta, int nctas, int epi_tid, uint32_t stg, uint64_t* wbar) { st.next = cta; }

    template <typename OutT>
    __device__ static void reduce_tile(const RankParams& p, int lm, int n_blk, int epi_tid) {
        const ccmpi_dev_comm& c = p.comm;
        if (epi_tid < c.size)
            ptx::wait_flag_ge_sys(user_flag(c, c.rank, flag_idx(p, lm, n_blk, epi_tid)), p.epoch, 620 + epi_tid);
        epi_bar_sync();
        constexpr int EPV = 16 / sizeof(OutT);
        const int BN = p.aux0, VPR = BN / EPV;
        const int rows = min(BM, p.chunk - lm * BM);
        const OutT* slots = reinterpret_cast<const OutT*>(c.heap[c.rank] + p.off1);
        for (int v = epi_tid; v < rows * VPR; v += EPI_THREADS) {
            const int lrow = lm * BM + v / VPR;
            const int col = n_blk * BN + (v % VPR) * EPV;
            float acc[EPV];
#pragma unroll
            for (int s = 0; s < CCMPI_MAX_RANKS; ++s) {
                if (s < c.size) {
                    const uint4 raw = ptx::ld_v4(slots + ((size_t)s * p.chunk + lrow) * p.N + col);
                    if constexpr (sizeof(OutT) == 2) {
                        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = __bfloat1622float2(h[e]);
                            acc[2 * e] = s == 0 ? f.x : acc[2 * e] + f.x;
                            acc[2 * e + 1] = s == 0 ? f.y : acc[2 * e + 1] + f.y;
                        }
                    } else {
                        const float* f = reinterpret_cast<const float*>(&raw);
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[e] = s == 0 ? f[e] : acc[e] + f[e];
                    }
                }
            }
            uint4 out;
            if constexpr (sizeof(OutT) == 2) {
                out.x = pack_bf16(__float_as_uint(acc[0]), __float_as_uint(acc[1]));
                out.y = pack_bf16(__float_as_uint(acc[2]), __float_as_uint(acc[3]));
                out.z = pack_bf16(__float_as_uint(acc[4]), __float_as_uint(acc[5]));
                out.w = pack_bf16(__float_as_uint(acc[6]), __float_as_uint(acc[7]));
            } else {
                out = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
            }
            ptx::st_v4(reinterpret_cast<OutT*>(p.C) + (size_t)lrow * p.ldc + col, out);
        }
    }

    template <typename OutT>
    __device__ static bool background(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        const int items = mbpr(p) * nblocks(p);
        if (st.next >= items) return false;
        const TileCoord tc = raster(st.next, mbpr(p), nblocks(p));
        const ccmpi_dev_comm& c = p.comm;
        if (epi_tid == 0) {
            uint32_t ok = 1;
            for (int s = 0; s < c.size; ++s)
                ok &= (int32_t)(ptx::ld_acquire_sys(user_flag(c, c.rank, flag_idx(p, tc.m_blk, tc.n_blk, s))) - p.epoch) >= 0 ? 1u : 0u;
            *bcast = ok;
        }
        epi_bar_sync();
        const bool ready = *bcast != 0;
        epi_bar_sync();
        if (!ready) return false;
        reduce_tile<OutT>(p, tc.m_blk, tc.n_blk, epi_tid);
        st.next += nctas;
        return true;
    }

    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        const int items = mbpr(p) * nblocks(p);
        for (; st.next < items; st.next += nctas) {
            const TileCoord tc = raster(st.next, mbpr(p), nblocks(p));
            reduce_tile<OutT>(p, tc.m_blk, tc.n_blk, epi_tid);
        }
    }
};

}  // namespace gemm

namespace gemm {

// ================================================================================================
// all-gather -> GEMM -> all-gather of the output: the reference's whole "naive" forward step
// (func_impl.py:76-109: gather the input, multiply, gather the output) in ONE kernel. The K-gather
// is AgKPolicy's background push; the epilogue is AgOutPolicy's: every finished tile is stored to
// column block `rank` of every member's y_full.
//   off0 / off1 / chunk / aux0 / aux2 : as in AgKPolicy (x_local, x_full, K/n, ldx, CTAs per rank)
//   off2 : symmetric offset of y_full [M][N * size] (OutT)
// The staging row of an epilogue lane serves both its pushes and its output rows; each use starts
// by waiting until the previous bulk copy has finished reading the row.
// Run on B200s in round 2 (tests/test_gpu_seq.py::test_gather_gemm_gather*). CTA-pair kernel only.
// ================================================================================================
struct AgKAgOutPolicy : AgKPolicy {
    static constexpr bool kEpiOps = false;   // own store: no fused elementwise epilogue
    static constexpr bool kStagedStore = true;

    template <typename OutT>
    __device__ static void bulk_row(const RankParams& p, const TileCoord& tc, int row, int col0, uint32_t smem_row, int bytes) {
        const ccmpi_dev_comm& c = p.comm;
        const size_t full_ld = (size_t)p.N * c.size;
        const size_t d = p.off2 + ((size_t)row * full_ld + (size_t)c.rank * p.N + col0) * sizeof(OutT);
#pragma unroll
        for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
            if (s < c.size) ptx::bulk_s2g(c.heap[(c.rank + s) % c.size] + d, smem_row, bytes);
    }

    // peers store into our y_full: they must not start before we are done with its previous contents
    __device__ static void prologue(const RankParams& p, int cta, int nctas, int epi_tid) {
        cta_rank_barrier(p.comm, cta, epi_tid, [] { epi_bar_sync(); });
    }

    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        AgKPolicy::tail<OutT>(p, st, cta, nctas, epi_tid, bcast);     // whatever has not been pushed yet
        if (epi_tid == 0) ptx::fence_acq_rel_sys();
        cta_rank_barrier(p.comm, cta, epi_tid, [] { epi_bar_sync(); });   // every tile has landed everywhere
    }
};

}  // namespace gemm
