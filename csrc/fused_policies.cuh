// Policies that fuse a collective into the persistent tcgen05 GEMM (gemm_kernel.cuh).
//
//   AgKPolicy   all-gather -> GEMM, gather axis = GEMM K   (func_impl.py:76-91 + the FC GEMM)
//   RsNPolicy   GEMM -> reduce-scatter, scatter axis = GEMM N (func_impl.py:150-187 after dx GEMM)
//   AgOutPolicy GEMM -> all-gather of the output on N       (func_impl.py:94-109 after the FC GEMM)
//   ArPolicy    GEMM -> all-reduce (row-parallel fc_o / fc2 forward: RsN + broadcast of the result)
//
// All cross-GPU traffic is issued from inside the kernel on peer-mapped pointers
// (ccmpi_dev_comm::heap[peer]) or on the NVLS multicast alias (mc_heap). Flags live in the
// user area of the communicator's signal pad and carry the launch epoch, so they never need
// resetting; buffers that peers write into are double-buffered by the caller (epoch parity),
// which is sufficient because a rank cannot get two launches ahead of a peer whose data it
// needs.
#pragma once
#include "gemm_kernel.cuh"

namespace gemm {

__device__ __forceinline__ uint32_t* user_flag(const ccmpi_dev_comm& c, int rank, int idx) {
    return c.pad[rank] + CCMPI_PAD_USER + idx;
}

// Cross-rank barrier for CTA `cta` of every rank, callable by any subset of threads of the CTA
// that contains thread `t0..t0+size-1` (uses the same pad words as the stand-alone collectives,
// csrc/collectives.cu: block_barrier). `sync` is the caller's intra-CTA barrier for the
// participating threads.
template <typename SyncFn>
__device__ __forceinline__ void cta_rank_barrier(const ccmpi_dev_comm& c, int cta, int tid, SyncFn sync) {
    sync();
    uint32_t* self = c.pad[c.rank];
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(self + CCMPI_PAD_EPOCH + cta) + 1;
    if (tid < c.size) {
        ptx::st_release_sys(c.pad[tid] + CCMPI_PAD_BARRIER + cta * CCMPI_MAX_RANKS + c.rank, epoch);
        ptx::wait_flag_ge_sys(self + CCMPI_PAD_BARRIER + cta * CCMPI_MAX_RANKS + tid, epoch, 900 + tid);
    }
    sync();
    if (tid == 0) *reinterpret_cast<volatile uint32_t*>(self + CCMPI_PAD_EPOCH + cta) = epoch;
}

// ================================================================================================
// all-gather -> GEMM, K-gather.   A = [x_0 | x_1 | ... | x_{n-1}] along K, x_r lives on rank r.
//   off0 : symmetric offset of x_local  [M][chunk]   (row stride aux0 elements)
//   off1 : symmetric offset of x_full   [M][K]       (staging the peers push into)
//   tmap_a  -> this rank's x_full,  tmap_a2 -> this rank's x_local
//   aux2 : CTAs per rank of this launch (contributions per ready flag and launch)
// The GEMM walks K starting with the local chunk (no wait), then chunks (rank+1)%n, ... each
// gated on ready[src][m_blk], which counts the pushing CTAs of rank `src` (see prologue()).
// ================================================================================================
struct AgKPolicy : PlainPolicy {
    static constexpr bool kCommCtas = false;

    __device__ static int ready_idx(const RankParams& p, int src, int m_blk) {
        return p.flag_base + src * ((p.M + BM - 1) / BM) + m_blk;
    }

    __device__ static int kcoord(const RankParams& p, int it) {
        const int per = p.chunk / BK;
        const int ci = it / per;
        return ((p.comm.rank + ci) % p.comm.size) * p.chunk + (it - ci * per) * BK;
    }

    __device__ static void kblock(const RankParams& p, const TileCoord& tc, int it, int& k0, const CUtensorMap*& ta,
                                  int& a_k0, int& a_m0) {
        const ccmpi_dev_comm& c = p.comm;
        const int per = p.chunk / BK;              // K-blocks per chunk
        const int ci = it / per, within = it - ci * per;
        const int src = (c.rank + ci) % c.size;
        k0 = src * p.chunk + within * BK;
        a_m0 = tc.m_blk * BM;
        if (src == c.rank) {
            ta = &p.tmap_a2;
            a_k0 = within * BK;
        } else {
            if (within == 0) {
                ptx::wait_flag_ge_sys(user_flag(c, c.rank, ready_idx(p, src, tc.m_blk)), p.epoch * (unsigned)p.aux2, 500 + src);
                ptx::fence_proxy_async_all();  // peer's generic-proxy stores -> our TMA reads
            }
            ta = &p.tmap_a;
            a_k0 = k0;
        }
    }

    __device__ static int num_kchunks(const RankParams& p) { return p.comm.size; }
    __device__ static void kchunk(const RankParams& p, const TileCoord& tc, int ci, bool wait, int& k_base,
                                  const CUtensorMap*& ta, int& a_k_base, int& a_m0) {
        const ccmpi_dev_comm& c = p.comm;
        int src = c.rank + ci;
        if (src >= c.size) src -= c.size;
        k_base = src * p.chunk;
        a_m0 = tc.m_blk * BM;
        if (src == c.rank) {
            ta = &p.tmap_a2;
            a_k_base = 0;
        } else {
            if (wait) {
                ptx::wait_flag_ge_sys(user_flag(c, c.rank, ready_idx(p, src, tc.m_blk)), p.epoch * (unsigned)p.aux2, 500 + src);
                ptx::fence_proxy_async_all();  // peer's stores -> our TMA reads
            }
            ta = &p.tmap_a;
            a_k_base = k_base;
        }
    }

    // The gather itself. There are no dedicated communication CTAs: the four epilogue warps of
    // EVERY CTA are idle whenever no accumulator is ready (always so at kernel start), and in
    // those gaps -- never while a finished accumulator is waiting -- they push this rank's shard -- 16-byte vectors, 8 loads in flight per thread, ~19 k threads
    // chip-wide -- into every member's x_full (their own included, which leaves the gathered
    // activation whole for the backward dW GEMM), in ascending m-block order, i.e. in the order
    // the peers' tile rasters will ask for it. After each unrolled batch a CTA adds 1 to
    // ready[rank][m_blk] on every peer for each m-block it has finished its share of; a flag is
    // complete when all `aux2` (= CTAs per rank) contributions of this launch are in. Meanwhile
    // the producer / MMA warps already work on the local K-chunk.
    static constexpr bool kBackground = true;
    static constexpr bool kPushFirst = true;
    static constexpr bool kNeedsStaging = true;
    struct State {
        long long base;   // first work-item index of this CTA's next batch (register path: 16-byte vectors)
        long long item;   // ... bulk path: first piece index of this CTA's next slice
        int pending;      // ... bulk path: m-blocks covered by the slices issued so far (published one slice later)
        bool local_started, local_done;   // ... bulk path: pass 1 (copy into this rank's own x_full)
        long long sf_item;                // ... bulk path, block-scaled operands: next scale run of this CTA
        bool sf_done;
        int flagged;      // m-blocks [0, flagged) already signalled by this CTA
        int m_blocks;
        uint32_t stg;     // this lane's staging row in shared memory (0: register path)
        uint64_t* wbar;   // this warp's mbarrier for the bulk loads
        uint32_t phase;
    };
    __device__ static void init_state(State& st, const RankParams& p, int cta, int nctas, int epi_tid, uint32_t stg, uint64_t* wbar) {
        st.base = (long long)cta * EPI_THREADS;
        st.item = (long long)cta * 4;
        st.pending = 0;
        st.local_started = st.local_done = false;
        st.sf_item = (long long)cta * 4;
        st.sf_done = false;
        st.flagged = 0;
        st.m_blocks = p.comm.size == 1 ? 0 : (p.M + BM - 1) / BM;
        st.stg = stg;
        st.wbar = wbar;
        st.phase = 0;
        if (stg != 0) {
            if ((epi_tid & 31) == 0) {
                ptx::mbar_init(wbar, 32);
                ptx::fence_mbar_init();
            }
            __syncwarp();
        }
    }

    // Bulk-copy flavour of the push (CTA-pair kernel). The TMA engine of an SM retires small bulk copies at
    // a fixed rate of a few tens of nanoseconds each, whatever their size (the first version moved one
    // 256-byte piece per epilogue LANE and measured 12 us per 32 KB slice -- 384 copies -- with the device
    // trace, i.e. ~400 GB/s over the whole chip and a producer that waited 135 us for gathered chunks). So
    // the copies are made as large as the staging area allows: each of the four epilogue WARPS moves one
    // piece of up to 8 KB per slice -- whole shard rows when a row is at most 4 KB, a column segment of one
    // row otherwise -- with ONE bulk load (the shard is contiguous) and one bulk store per row and member,
    // issued by lane 0. Remote members are served first; the copy into this rank's own x_full (which only
    // the backward dW GEMM reads) goes last.
    static constexpr int PUSH_PIECE = 8192;
    template <typename OutT>
    __device__ static bool background(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        const ccmpi_dev_comm& c = p.comm;
        // Groups larger than two with an NVLS multicast mapping keep the register path: one
        // multimem.st per vector leaves the GPU once and is replicated by the switch, while the
        // bulk path would send n-1 unicast copies (measured at mp = 8: 5.55 vs 5.73 ms per step).
        if (st.stg == 0 || ((p.aux1 & 1) != 0 && c.mc_heap != nullptr && c.size > 2))
            return background_regs<OutT>(p, st, cta, nctas, epi_tid, bcast);
        // Two passes over the shard. Pass 0 serves the PEERS, who are waiting for it, and publishes the
        // ready flags; pass 1 copies the shard into this rank's own x_full, which nothing reads before the
        // backward dW GEMM -- so it must not share the copy engine's time with what the peers' tensor
        // cores are starved for.
        if (st.m_blocks == 0 || st.local_done) return false;
        const int lane_ = epi_tid & 31, w_ = epi_tid >> 5;
        if (p.sf_run_bytes > 0 && !st.sf_done) {
            // Block-scaled operands: this rank's scale chunks travel first -- one run of sf_run_bytes per m-block,
            // into every peer's gathered scale array at this rank's position in the K order. They are issued (and,
            // being earlier bulk groups, completed) before any payload piece of this CTA is published.
            int segs = 1;
            while (p.sf_run_bytes / segs > PUSH_PIECE) segs <<= 1;
            const int seg_bytes = p.sf_run_bytes / segs;
            const long long total = (long long)st.m_blocks * segs;
            const long long i = st.sf_item + w_;
            const bool ok = i < total;
            const uint32_t wbuf = st.stg - (uint32_t)lane_ * STG_ROW_STRIDE;
            if (lane_ == 0) ptx::tma_store_wait_read<0>();
            ptx::mbar_arrive_expect_tx(st.wbar, (lane_ == 0 && ok) ? (uint32_t)seg_bytes : 0u);
            const long long mb = ok ? i / segs : 0;
            const int seg = ok ? (int)(i - mb * segs) : 0;
            if (lane_ == 0 && ok) {
                const uint8_t* src = c.heap[c.rank] + p.off_sf_local + (size_t)mb * p.sf_run_bytes + (size_t)seg * seg_bytes;
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(wbuf),
                             "l"(src), "r"(seg_bytes), "r"(ptx::smem_u32(st.wbar))
                             : "memory");
            }
            ptx::mbar_wait(st.wbar, st.phase, 801);
            st.phase ^= 1;
            if (lane_ == 0 && ok) {
                const size_t d = p.off_sf_full + ((size_t)mb * c.size + c.rank) * p.sf_run_bytes + (size_t)seg * seg_bytes;
                for (int s2 = 1; s2 < c.size; ++s2) ptx::bulk_s2g(c.heap[(c.rank + s2) % c.size] + d, wbuf, seg_bytes);
            }
            ptx::tma_store_commit();
            st.sf_item += (long long)nctas * 4;
            if (st.sf_item >= total) st.sf_done = true;
            return true;
        }
        const bool local_pass = st.flagged >= st.m_blocks;
        if (local_pass && (p.aux1 & 8)) {     // the caller does not need the gathered operand afterwards
            st.local_done = true;
            return false;
        }
        if (local_pass && !st.local_started) {
            st.local_started = true;
            st.item = (long long)cta * 4;
        }
        const int lane = epi_tid & 31, w = epi_tid >> 5;
        const int row_bytes = p.chunk * 2;
        int segs = 1;                                              // column segments per row (power of two)
        while (row_bytes / segs > PUSH_PIECE) segs <<= 1;
        const int seg_bytes = row_bytes / segs;
        int rows_per = 1;                                          // whole rows per piece (power of two), contiguous source only
        if (segs == 1 && p.aux0 == p.chunk)
            while (rows_per * 2 * row_bytes <= PUSH_PIECE && rows_per < BM) rows_per <<= 1;
        const long long items_per_blk = (long long)(BM / rows_per) * segs;
        const long long total = ((long long)(p.M + rows_per - 1) / rows_per) * segs;
        const long long stride = (long long)nctas * 4;            // one piece per epilogue warp and slice
        const uint8_t* x = c.heap[c.rank] + p.off0;
        // (row strides in bytes; the gathered row is n shards wide -- not written as K * 2 because fp8 operands
        // count K in elements while the policy's chunk / ldx are in 2-byte units)
        const size_t src_ld = (size_t)p.aux0 * 2, dst_ld = (size_t)row_bytes * c.size;
        const size_t col0 = (size_t)c.rank * row_bytes;
        const long long i = st.item + w;
        const bool ok = i < total;
        const uint32_t wbuf = st.stg - (uint32_t)lane * STG_ROW_STRIDE;          // this warp's 8.5 KB of staging
        long long row0 = 0;
        int seg = 0, nrows = 0;
        if (ok) {
            const long long ri = i / segs;
            seg = (int)(i - ri * segs);
            row0 = ri * rows_per;
            nrows = (int)min((long long)rows_per, (long long)p.M - row0);
        }
        const uint32_t bytes = ok ? (uint32_t)(nrows * seg_bytes) : 0u;
        if (lane == 0) ptx::tma_store_wait_read<0>();              // my previous piece has left the staging buffer
        ptx::mbar_arrive_expect_tx(st.wbar, lane == 0 ? bytes : 0u);
        if (lane == 0 && ok) {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(wbuf),
                         "l"(x + (size_t)row0 * src_ld + (size_t)seg * seg_bytes), "r"(bytes), "r"(ptx::smem_u32(st.wbar))
                         : "memory");
        }
        ptx::mbar_wait(st.wbar, st.phase, 800);
        st.phase ^= 1;
        if (lane == 0 && ok) {
            const int s0 = local_pass ? 0 : 1, s1 = local_pass ? 1 : c.size;
#pragma unroll 1
            for (int s = s0; s < s1; ++s) {
                uint8_t* dst = c.heap[(c.rank + s) % c.size] + p.off1 + (size_t)row0 * dst_ld + col0 + (size_t)seg * seg_bytes;
                for (int r = 0; r < nrows; ++r) ptx::bulk_s2g(dst + (size_t)r * dst_ld, wbuf + (uint32_t)(r * seg_bytes), seg_bytes);
            }
        }
        ptx::tma_store_commit();
        if (local_pass) {
            st.item += stride;
            if (st.item >= total) {                        // this CTA has no pieces left
                ptx::tma_store_wait<0>();
                st.local_done = true;
            }
            return true;
        }
        // every piece of THIS CTA below `next` is on its way (its later pieces start at next + w)
        const long long next = st.item + stride;
        int done = (int)min((long long)st.m_blocks, next / items_per_blk);
        if (next >= total) done = st.m_blocks;
        st.item = next;
        // Publication runs ONE SLICE BEHIND the copies. Waiting for the slice just issued to land at the
        // peer costs an NVLink round trip under load (the device trace: ~13 us per slice, 90 us for a 33 MB
        // shard, whatever the copy size); the previous slice's copies have had a whole slice to complete,
        // so `wait_group 1` returns at once and the round trip overlaps the next load.
        // (The very first slice is published at once: until it is, no peer can start on its remote chunk anyway.)
        int publish;
        if (st.flagged == 0 && st.pending == 0) {
            ptx::tma_store_wait<0>();
            publish = done;
        } else {
            ptx::tma_store_wait<1>();      // every group but the newest has landed ...
            publish = st.pending;          // ... which is what `pending` describes
        }
        st.pending = done;
        if (publish <= st.flagged) return true;
        ptx::fence_proxy_async_all();      // landed copies are ordered before the flag traffic
        epi_bar_sync();
        if (epi_tid == 0) {
            ptx::fence_acq_rel_sys();
            for (int j = st.flagged; j < publish; ++j)
                for (int s = 1; s < c.size; ++s)
                    ptx::red_relaxed_sys_add(user_flag(c, (c.rank + s) % c.size, ready_idx(p, c.rank, j)), 1u);
        }
        st.flagged = publish;
        return true;
    }

    // Register flavour (single-CTA kernel, which has no staging rows): 8 vectors per thread.
    template <typename OutT>
    __device__ static bool background_regs(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        if (st.flagged >= st.m_blocks) return false;
        const ccmpi_dev_comm& c = p.comm;
        const int vpr = p.chunk / 8;                        // 16-byte vectors per shard row
        const long long total = (long long)p.M * vpr;
        const long long per_blk = (long long)BM * vpr;
        const long long stride = (long long)nctas * EPI_THREADS;
        const uint8_t* x = c.heap[c.rank] + p.off0;
        const size_t src_ld = (size_t)p.aux0 * 2, dst_ld = (size_t)p.K * 2;
        const size_t col0 = (size_t)c.rank * p.chunk * 2;
        const bool mc = (p.aux1 & 1) != 0 && c.mc_heap != nullptr;
        constexpr int U = 8;
        uint4 v[U];
        size_t d[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = st.base + epi_tid + u * stride;
            ok[u] = i < total;
            if (ok[u]) {
                const long long row = i / vpr;
                const int w = (int)(i - row * vpr);
                v[u] = ptx::ld_nc_v4(x + (size_t)row * src_ld + (size_t)w * 16);
                d[u] = p.off1 + (size_t)row * dst_ld + col0 + (size_t)w * 16;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (ok[u]) {
                if (mc) {
                    ptx::multimem_st_v4(c.mc_heap + d[u], v[u]);
                } else {
#pragma unroll
                    for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
                        if (s < c.size) ptx::st_v4(c.heap[(c.rank + s) % c.size] + d[u], v[u]);
                }
            }
        }
        // every vector of this CTA below `next` is on its way
        const long long next = st.base + U * stride;
        int done = (int)min((long long)st.m_blocks, next / per_blk);
        // Flags go out once per raster group (16 m-blocks = what the consumers' rasters sweep
        // before moving on), not once per batch: each signal costs a system-scope fence that
        // waits for this CTA's remote stores, and the other epilogue threads wait for it.
        constexpr int GRANULE = 16;
        done = (done / GRANULE) * GRANULE;
        if (next >= total) done = st.m_blocks;
        if (done <= st.flagged) {       // uniform across the CTA: nothing to publish yet
            st.base = next;
            return true;
        }
        epi_bar_sync();
        if (epi_tid == 0 && done > st.flagged) {
            // one system-scope fence orders every push of this CTA (cumulative over the barrier)
            // before the flag updates; the updates themselves can then be relaxed and pipelined
            ptx::fence_acq_rel_sys();
            for (int j = st.flagged; j < done; ++j)
                for (int s = 1; s < c.size; ++s)
                    ptx::red_relaxed_sys_add(user_flag(c, (c.rank + s) % c.size, ready_idx(p, c.rank, j)), 1u);
        }
        st.flagged = done;
        st.base = next;
        return true;
    }

    // whatever has not been pushed by the time this CTA runs out of tiles goes out now: peers
    // depend on it
    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        while (background<OutT>(p, st, cta, nctas, epi_tid, bcast)) {}
    }
};

// ================================================================================================
// GEMM -> reduce-scatter, N-scatter.  C_partial[M][N] is produced tile by tile; column block
// [r*chunk, (r+1)*chunk) belongs to rank r.
//   off1 : symmetric offset of the slot buffer  [n sources][M][chunk]  (OutT)
//   C    : this rank's reduced output            [M][chunk], row stride ldc
// The epilogue stores each partial tile into the OWNER's slot for this source rank and raises
// flag[tile][source] there. Tiles owned by peers are computed first. When a CTA runs out of
// tiles its epilogue warps turn into reducers: they wait for all n sources of an owned tile and
// sum the slots in rank order 0..n-1 (the order of np.sum(axis=0) in func_impl.py:186).
// ================================================================================================
//   aux0 : the launch's BN (tile width), so one policy type serves every tile shape
struct RsNSlotPolicy : PlainPolicy {
    using SingleCta = RsNSlotPolicy;
    static constexpr bool kEpiOps = false;   // own store: no fused elementwise epilogue
    static constexpr bool kCommCtas = false;
    static constexpr bool kStagedStore = true;
    static constexpr bool kNeedsStaging = true;

    // CTA-pair kernel: one row chunk of a partial tile, already in shared memory, goes to the
    // owner's slot for this source rank in a single bulk copy
    template <typename OutT>
    __device__ static void bulk_row(const RankParams& p, const TileCoord& tc, int row, int col0, uint32_t smem_row, int bytes) {
        const ccmpi_dev_comm& c = p.comm;
        const int owner = col0 / p.chunk;
        const int lcol = col0 - owner * p.chunk;
        OutT* slot = reinterpret_cast<OutT*>(c.heap[owner] + p.off1) + ((size_t)c.rank * p.M + row) * p.chunk + lcol;
        ptx::bulk_s2g(slot, smem_row, bytes);
    }

    __device__ static int nb_per_owner(const RankParams& p) { return p.chunk / p.aux0; }
    __device__ static int flag_idx(const RankParams& p, int m_blk, int n_local, int src) {
        return p.flag_base + ((m_blk * nb_per_owner(p) + n_local) * p.comm.size + src);
    }

    // Owners are interleaved tile by tile: rank+1, rank+2, ..., rank, rank+1, ... That spreads the
    // remote stores over the whole kernel instead of bursting them at the start, and -- every
    // rank using the mirrored order -- a peer's partial for one of our tiles arrives at about the
    // time we finish the same tile ourselves, ready for the background reducer.
    __device__ static TileCoord tile(const RankParams& p, int t, int mb, int nb) {
        const int n = p.comm.size;
        const int g = t % n, in_g = t / n;
        const int owner = (p.comm.rank + 1 + g) % n;
        TileCoord tc = raster(in_g, mb, nb_per_owner(p));
        tc.n_blk += owner * nb_per_owner(p);
        return tc;
    }

    template <typename OutT>
    __device__ static void store(const RankParams& p, const TileCoord& tc, int row, int col0, const uint32_t (&v)[32]) {
        if (row >= p.M) return;
        const ccmpi_dev_comm& c = p.comm;
        const int owner = col0 / p.chunk;
        const int lcol = col0 - owner * p.chunk;
        OutT* slot = reinterpret_cast<OutT*>(c.heap[owner] + p.off1) +
                     ((size_t)c.rank * p.M + row) * p.chunk + lcol;
        store_row32<OutT>(slot, v, 32, false);
    }

    __device__ static void tile_done(const RankParams& p, const TileCoord& tc, int epi_tid) {
        epi_bar_sync();
        if (epi_tid == 0) {
            const ccmpi_dev_comm& c = p.comm;
            const int owner = tc.n_blk / nb_per_owner(p);
            const int n_local = tc.n_blk - owner * nb_per_owner(p);
            ptx::fence_acq_rel_sys();
            ptx::st_release_sys(user_flag(c, owner, flag_idx(p, tc.m_blk, n_local, c.rank)), p.epoch);
        }
    }

    // Reduce one owned tile: 128 threads, 16-byte vectors, rank-ordered fp32 accumulation.
    template <typename OutT, bool BROADCAST>
    __device__ static void reduce_tile(const RankParams& p, int m_blk, int n_local, int epi_tid) {
        const ccmpi_dev_comm& c = p.comm;
        if (epi_tid < c.size)
            ptx::wait_flag_ge_sys(user_flag(c, c.rank, flag_idx(p, m_blk, n_local, epi_tid)), p.epoch, 600 + epi_tid);
        epi_bar_sync();
        constexpr int EPV = 16 / sizeof(OutT);          // elements per 16-byte vector
        const int BN = p.aux0;
        const int VPR = BN / EPV;                       // vectors per tile row
        const int rows = min(BM, p.M - m_blk * BM);
        const OutT* slots = reinterpret_cast<const OutT*>(c.heap[c.rank] + p.off1);
        constexpr int U = 4;   // vectors per thread in flight (x n sources): the slots sit in HBM / far L2
        // EPI_ADD (bf16 reduce-scatter only): the residual / incoming gradient [M][chunk] (row stride
        // epi_ld) is added to the reduced tile on its way out, in fp32, before the one rounding
        const bool add_res = !BROADCAST && sizeof(OutT) == 2 && p.epi_op == EPI_ADD;
        for (int v0 = epi_tid; v0 < rows * VPR; v0 += U * EPI_THREADS) {
            uint4 raw[U][CCMPI_MAX_RANKS];
            uint4 res[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int v = v0 + u * EPI_THREADS;
                if (v < rows * VPR) {
                    const int row = m_blk * BM + v / VPR;
                    const int lcol = n_local * BN + (v % VPR) * EPV;
#pragma unroll
                    for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
                        if (s < c.size) raw[u][s] = ptx::ld_v4(slots + ((size_t)s * p.M + row) * p.chunk + lcol);
                    if (add_res)
                        res[u] = ptx::ld_v4(reinterpret_cast<const __nv_bfloat16*>(p.epi_in) + (size_t)row * p.epi_ld + lcol);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int v = v0 + u * EPI_THREADS;
                if (v >= rows * VPR) continue;
                const int row = m_blk * BM + v / VPR;
                const int lcol = n_local * BN + (v % VPR) * EPV;
                float acc[EPV];
#pragma unroll
                for (int s = 0; s < CCMPI_MAX_RANKS; ++s) {
                    if (s < c.size) {
                        if constexpr (sizeof(OutT) == 2) {
                            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[u][s]);
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 f = __bfloat1622float2(h[e]);
                                acc[2 * e] = s == 0 ? f.x : acc[2 * e] + f.x;
                                acc[2 * e + 1] = s == 0 ? f.y : acc[2 * e + 1] + f.y;
                            }
                        } else {
                            const float* f = reinterpret_cast<const float*>(&raw[u][s]);
#pragma unroll
                            for (int e = 0; e < 4; ++e) acc[e] = s == 0 ? f[e] : acc[e] + f[e];
                        }
                    }
                }
                uint4 out;
                if constexpr (sizeof(OutT) == 2) {
                    if (add_res) {
                        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&res[u]);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = __bfloat1622float2(h[e]);
                            acc[2 * e] += f.x;
                            acc[2 * e + 1] += f.y;
                        }
                    }
                    out.x = pack_bf16(__float_as_uint(acc[0]), __float_as_uint(acc[1]));
                    out.y = pack_bf16(__float_as_uint(acc[2]), __float_as_uint(acc[3]));
                    out.z = pack_bf16(__float_as_uint(acc[4]), __float_as_uint(acc[5]));
                    out.w = pack_bf16(__float_as_uint(acc[6]), __float_as_uint(acc[7]));
                } else {
                    out = make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]),
                                     __float_as_uint(acc[3]));
                }
                if constexpr (!BROADCAST) {
                    ptx::st_v4(reinterpret_cast<OutT*>(p.C) + (size_t)row * p.ldc + lcol, out);
                } else {
                    // all-reduce flavour: the reduced tile goes to every rank's full-width output
                    // (off2: symmetric [M][N], row stride N)
                    const size_t d = p.off2 + ((size_t)row * p.N + (size_t)c.rank * p.chunk + lcol) * sizeof(OutT);
                    if ((p.aux1 & 1) != 0 && c.mc_heap != nullptr) {
                        ptx::multimem_st_v4(c.mc_heap + d, out);
                    } else {
#pragma unroll
                        for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
                            if (s < c.size) ptx::st_v4(c.heap[(c.rank + s) % c.size] + d, out);
                    }
                }
            }
        }
    }

    // Reducing owned tiles is background work too: between two of its own tiles a CTA checks
    // whether every source of its next owned tile has arrived and, if so, reduces it right away,
    // so that little is left for the tail.
    static constexpr bool kBackground = true;
    struct State { int next; };
    __device__ static void init_state(State& st, const RankParams& p, int cta, int nctas, int epi_tid, uint32_t stg, uint64_t* wbar) { st.next = cta; }

    template <typename OutT, bool BROADCAST>
    __device__ static bool try_reduce(const RankParams& p, State& st, int nctas, int epi_tid, volatile uint32_t* bcast) {
        const int mb = (p.M + BM - 1) / BM, per = nb_per_owner(p);
        if (st.next >= mb * per) return false;
        const TileCoord tc = raster(st.next, mb, per);
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
        reduce_tile<OutT, BROADCAST>(p, tc.m_blk, tc.n_blk, epi_tid);   // its flag wait returns at once
        st.next += nctas;
        return true;
    }
    template <typename OutT>
    __device__ static bool background(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        return try_reduce<OutT, false>(p, st, nctas, epi_tid, bcast);
    }
    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        const int mb = (p.M + BM - 1) / BM, per = nb_per_owner(p);
        for (; st.next < mb * per; st.next += nctas) {
            const TileCoord tc = raster(st.next, mb, per);
            reduce_tile<OutT, false>(p, tc.m_blk, tc.n_blk, epi_tid);
        }
    }
};

// ------------------------------------------------------------------------------------------------
// GEMM -> reduce-scatter for the CTA-pair kernel: OWN TILES LAST, reduced in their own epilogue.
//
// RsNSlotPolicy writes every partial tile -- the owner's own included -- to a slot and lets reducer
// passes sum the slots afterwards; the device trace showed those passes (latency-bound 128-thread loops
// in the gaps and in the tail) costing 70-130 us per kernel at mp = 2. Here a rank first computes the
// tiles its PEERS own (interleaved over the peers, pushed into their slots exactly as before) and its
// own tiles at the end. By then the peers, who run the mirrored order, have long delivered their
// partials for those tiles, so the epilogue of an own tile adds them to its accumulator straight out of
// TMEM -- rank-ordered, every partial rounded to the slot type first, i.e. bit-identical to the slot
// path and to np.sum(axis=0) order (func_impl.py:186) -- plus the optional residual (EPI_ADD), and
// stores the final values. No slot write for own tiles, no reducer, no tail.
// ------------------------------------------------------------------------------------------------
struct RsNPolicy : RsNSlotPolicy {
    using SingleCta = RsNSlotPolicy;        // the single-CTA kernel keeps the slot path
    static constexpr bool kOwnTileDirect = true;
    static constexpr bool kBackground = false;
    struct State {};
    __device__ static void init_state(State&, const RankParams&, int, int, int, uint32_t, uint64_t*) {}
    template <typename OutT>
    __device__ static bool background(const RankParams&, State&, int, int, int, volatile uint32_t*) { return false; }
    template <typename OutT>
    __device__ static void tail(const RankParams&, State&, int, int, int, volatile uint32_t*) {}

    __device__ static TileCoord tile(const RankParams& p, int t, int mb, int nb) {
        const int n = p.comm.size, per = nb_per_owner(p);
        const int remote = (n - 1) * mb * per;
        int owner, in_g;
        if (t < remote) {
            const int g = t % (n - 1);
            in_g = t / (n - 1);
            owner = (p.comm.rank + 1 + g) % n;
        } else {
            owner = p.comm.rank;
            in_g = t - remote;
        }
        TileCoord tc = raster(in_g, mb, per);
        tc.n_blk += owner * per;
        return tc;
    }
    __device__ static bool is_own(const RankParams& p, const TileCoord& tc) { return tc.n_blk / nb_per_owner(p) == p.comm.rank; }

    __device__ static void tile_done(const RankParams& p, const TileCoord& tc, int epi_tid) {
        if (!is_own(p, tc)) RsNSlotPolicy::tile_done(p, tc, epi_tid);
    }
    // all 128 epilogue threads: every peer's partial of this own tile has landed in its slot
    __device__ static void own_wait(const RankParams& p, const TileCoord& tc, int epi_tid) {
        const ccmpi_dev_comm& c = p.comm;
        const int n_local = tc.n_blk - c.rank * nb_per_owner(p);
        if (epi_tid < c.size && epi_tid != c.rank)
            ptx::wait_flag_ge_sys(user_flag(c, c.rank, flag_idx(p, tc.m_blk, n_local, epi_tid)), p.epoch, 600 + epi_tid);
        epi_bar_sync();
    }

    // What an own tile adds to its accumulator, as a list of "pseudo-sources": the peers' slots in rank
    // order, then the residual. The first PF of them are requested one 32-column chunk ahead of their use.
    static constexpr int PF = 2;
    template <typename OutT>
    struct OwnPf {
        uint4 v[PF][32 * sizeof(OutT) / 16];
    };
    __device__ static int own_nsrc(const RankParams& p) { return p.comm.size - 1 + (p.epi_op == EPI_ADD ? 1 : 0); }
    template <typename OutT>
    __device__ static void own_load(const RankParams& p, int k, int row, int lcol, uint4 (&out)[32 * sizeof(OutT) / 16]) {
        constexpr int VPC = 32 * sizeof(OutT) / 16;
        const ccmpi_dev_comm& c = p.comm;
        const uint8_t* src;
        if (k < c.size - 1) {
            const int s = k < c.rank ? k : k + 1;
            src = c.heap[c.rank] + p.off1 + (((size_t)s * p.M + row) * p.chunk + lcol) * sizeof(OutT);
        } else {
            src = reinterpret_cast<const uint8_t*>(p.epi_in) + ((size_t)row * p.epi_ld + lcol) * 2;   // bf16 residual
        }
        const int nv = (k < c.size - 1) ? VPC : 4;
#pragma unroll
        for (int j = 0; j < VPC; ++j)
            if (j < nv) out[j] = ptx::ld_v4(src + 16 * j);
    }
    template <typename OutT>
    __device__ static void own_prefetch(const RankParams& p, int row, int col0, OwnPf<OutT>& pf) {
        if (row >= p.M) return;
        const int lcol = col0 - p.comm.rank * p.chunk;
        const int ns = own_nsrc(p);
#pragma unroll
        for (int k = 0; k < PF; ++k)
            if (k < ns) own_load<OutT>(p, k, row, lcol, pf.v[k]);
    }
    template <typename OutT>
    __device__ static void own_store(const RankParams& p, const TileCoord& tc, int row, int col0, uint32_t (&v)[32],
                                     const OwnPf<OutT>& pf) {
        if (row >= p.M) return;
        constexpr int VPC = 32 * sizeof(OutT) / 16;
        const ccmpi_dev_comm& c = p.comm;
        const int lcol = col0 - c.rank * p.chunk;
        const int ns = own_nsrc(p);
        float own[32], acc[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            float f = __uint_as_float(v[i]);
            if constexpr (sizeof(OutT) == 2) f = __bfloat162float(__float2bfloat16_rn(f));   // what a slot would hold
            own[i] = f;
        }
        // Rank-ordered sum with this rank's partial at position `rank`. Every index into the prefetch
        // registers is a compile-time constant (a run-time index would push them to local memory).
        bool started = false;
        auto add_own = [&] {
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[i] = started ? acc[i] + own[i] : own[i];
            started = true;
        };
        auto add_bf16 = [&](const uint4* src) {
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(src);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 f = __bfloat1622float2(h[i]);
                acc[2 * i] = started ? acc[2 * i] + f.x : f.x;
                acc[2 * i + 1] = started ? acc[2 * i + 1] + f.y : f.y;
            }
            started = true;
        };
        auto add_slot = [&](const uint4* src) {
            if constexpr (sizeof(OutT) == 2) {
                add_bf16(src);
            } else {
                const float* f = reinterpret_cast<const float*>(src);
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = started ? acc[i] + f[i] : f[i];
                started = true;
            }
        };
#pragma unroll
        for (int k = 0; k < CCMPI_MAX_RANKS - 1; ++k) {
            if (k < c.size - 1) {
                if (k == c.rank) add_own();
                if (k < PF) {
                    add_slot(pf.v[k < PF ? k : 0]);
                } else {                            // more peers than prefetch slots: fetch now
                    uint4 late[VPC];
                    own_load<OutT>(p, k, row, lcol, late);
                    add_slot(late);
                }
            }
        }
        if (c.rank >= c.size - 1) add_own();
        if (p.epi_op == EPI_ADD) {
            const int kr = c.size - 1;              // the residual is the last pseudo-source
            bool done = false;
#pragma unroll
            for (int k = 0; k < PF; ++k)
                if (k == kr) {
                    add_bf16(pf.v[k]);
                    done = true;
                }
            if (!done) {
                uint4 late[VPC];
                own_load<OutT>(p, kr, row, lcol, late);
                add_bf16(late);
            }
        }
        (void)ns;
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(acc[i]);
        store_row32<OutT>(reinterpret_cast<OutT*>(p.C) + (size_t)row * p.ldc + lcol, v, 32, false);
    }
};

// GEMM -> all-reduce: RsN, but the reducer broadcasts and a final cross-rank barrier makes sure
// every rank's output is complete when its kernel ends.
struct ArPolicy : RsNSlotPolicy {
    using SingleCta = ArPolicy;
    // peers' kernels store into our output: they must not start before we are done with the
    // previous contents
    __device__ static void prologue(const RankParams& p, int cta, int nctas, int epi_tid) {
        cta_rank_barrier(p.comm, cta, epi_tid, [] { epi_bar_sync(); });
    }
    template <typename OutT>
    __device__ static bool background(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        return try_reduce<OutT, true>(p, st, nctas, epi_tid, bcast);
    }
    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        const int mb = (p.M + BM - 1) / BM, per = nb_per_owner(p);
        for (; st.next < mb * per; st.next += nctas) {
            const TileCoord tc = raster(st.next, mb, per);
            reduce_tile<OutT, true>(p, tc.m_blk, tc.n_blk, epi_tid);
        }
        if (epi_tid == 0) ptx::fence_acq_rel_sys();
        cta_rank_barrier(p.comm, cta, epi_tid, [] { epi_bar_sync(); });
    }
};

// ================================================================================================
// GEMM -> all-gather of the output along N.  y_local[M][N] is this rank's column block of
// y_full[M][N*n]; every finished tile is stored straight into every rank's y_full.
//   off1 : symmetric offset of y_full  [M][N * size]  (OutT)
// A cross-rank barrier before the first remote store (peers have finished reading the previous
// y_full) and one at the end (all tiles have landed everywhere) bracket the stores; the first
// overlaps the first tile's mainloop.
// ================================================================================================
struct AgOutPolicy : PlainPolicy {
    static constexpr bool kEpiOps = false;   // own store: no fused elementwise epilogue
    static constexpr bool kStagedStore = true;
    static constexpr bool kNeedsStaging = true;

    template <typename OutT>
    __device__ static void bulk_row(const RankParams& p, const TileCoord& tc, int row, int col0, uint32_t smem_row, int bytes) {
        const ccmpi_dev_comm& c = p.comm;
        const size_t full_ld = (size_t)p.N * c.size;
        const size_t d = p.off1 + ((size_t)row * full_ld + (size_t)c.rank * p.N + col0) * sizeof(OutT);
#pragma unroll
        for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
            if (s < c.size) ptx::bulk_s2g(c.heap[(c.rank + s) % c.size] + d, smem_row, bytes);
    }

    __device__ static void prologue(const RankParams& p, int cta, int nctas, int epi_tid) {
        cta_rank_barrier(p.comm, cta, epi_tid, [] { epi_bar_sync(); });
    }
    template <typename OutT>
    __device__ static void tail(const RankParams& p, State& st, int cta, int nctas, int epi_tid, volatile uint32_t* bcast) {
        if (epi_tid == 0) ptx::fence_acq_rel_sys();
        cta_rank_barrier(p.comm, cta, epi_tid, [] { epi_bar_sync(); });
    }
    template <typename OutT>
    __device__ static void store(const RankParams& p, const TileCoord& tc, int row, int col0, const uint32_t (&v)[32]) {
        if (row >= p.M) return;
        const ccmpi_dev_comm& c = p.comm;
        const int ncols = min(32, p.N - col0);
        if (ncols <= 0) return;
        const size_t full_ld = (size_t)p.N * c.size;
        const size_t d = p.off1 + ((size_t)row * full_ld + (size_t)c.rank * p.N + col0) * sizeof(OutT);
        if ((p.aux1 & 1) != 0 && c.mc_heap != nullptr && ncols == 32) {
            // one multimem.st per 16 bytes; the switch replicates it to every member
            if constexpr (sizeof(OutT) == 2) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint4 q;
                    q.x = pack_bf16(v[8 * j + 0], v[8 * j + 1]);
                    q.y = pack_bf16(v[8 * j + 2], v[8 * j + 3]);
                    q.z = pack_bf16(v[8 * j + 4], v[8 * j + 5]);
                    q.w = pack_bf16(v[8 * j + 6], v[8 * j + 7]);
                    ptx::multimem_st_v4(c.mc_heap + d + 16 * j, q);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    ptx::multimem_st_v4(c.mc_heap + d + 16 * j, make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
            }
        } else {
#pragma unroll
            for (int s = 0; s < CCMPI_MAX_RANKS; ++s)
                if (s < c.size) store_row32<OutT>(reinterpret_cast<OutT*>(c.heap[(c.rank + s) % c.size] + d), v, ncols, false);
        }
    }
};

}  // namespace gemm
