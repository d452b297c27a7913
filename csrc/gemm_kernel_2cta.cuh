// CTA-pair (cta_group::2) variant of the persistent tcgen05 GEMM kernel.
//
// Two CTAs of a 2-CTA cluster (the two SMs of a TPC) compute one 256 x BN tile: each loads ITS
// 128 rows of A and ITS half of the B tile, the leader's elected thread issues
// tcgen05.mma.cta_group::2 (M = 256) which reads both CTAs' shared memory and writes each CTA's
// 128 accumulator rows into that CTA's own TMEM. Per unit of MMA work a CTA pulls 2/3 of the
// bytes the single-CTA kernel needs (L2 -> SM traffic is what bounds the 128 x 256 single-CTA
// tile on B200), and the shared-memory read rate per SM is halved.
//
// From a policy's point of view nothing changes: every CTA still owns a 128-row m-block and a
// full BN-wide n-block -- CTA r of a pair gets m_blk = 2*pair_m + r -- so the fused-collective
// policies (gather flags per m-block, per-tile scatter flags, per-CTA rank barriers) are reused
// as they are.
//
// Barrier topology (per stage s / accumulator a):
//   full[s]        lives in the leader; 2 arrivals (leader's expect_tx + peer's remote arrive),
//                  transaction bytes from BOTH CTAs' TMA loads (their mbarrier operand is masked
//                  to the leader's copy)
//   empty[s]       one per CTA; tcgen05.commit.multicast arrives on both when the MMAs retire
//   tmem_full[a]   one per CTA; multicast commit
//   tmem_empty[a]  lives in the leader; 8 arrivals (4 epilogue warps of each CTA)
#pragma once
#include "gemm_kernel.cuh"
#include "mx_desc.cuh"

namespace gemm {

// Staged epilogue (policies whose tiles leave the GPU): every epilogue lane parks 256 bytes of its
// output row in shared memory and ships them with one cp.async.bulk per destination, so NVLink
// sees 256-byte bursts instead of 16-byte stores from registers. Rows are padded by 16 bytes,
// which makes the lanes' 16-byte shared stores conflict-free.
// (STG_CHUNK_BYTES / STG_ROW_STRIDE / STG_BYTES live in gemm_tcgen05.cuh: the policies use them too)

template <int BN, bool STAGED = false, bool MX = false>
struct Cfg2 {
    static constexpr int A_BYTES = BM * BK * 2;          // this CTA's 128 rows of A
    static constexpr int B_ROWS = BN / 2;                // this CTA's half of the B tile
    static constexpr int B_BYTES = B_ROWS * BK * 2;
    // block-scaled (MXFP8) operands: every stage also carries one 512-byte chunk of scales for this CTA's 128 rows
    // of A and one for the tile's 128 columns of B (see gemm_mxfp8.cu / DESIGN.md section 6)
    static constexpr int SF_BYTES = MX ? 512 : 0;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + 2 * SF_BYTES;
    static constexpr int SF_COL0 = 2 * BN;               // first scale column of tensor memory (MX)
    static_assert(!MX || BN == 128, "block-scaled tiles are 256 x 128 per CTA pair");
    static_assert(STAGE_BYTES % 1024 == 0, "operand tiles must stay 1024-byte aligned");
    static constexpr int RING_BUDGET = 200 * 1024 - (STAGED ? STG_BYTES : 0);
    static constexpr int STAGES = RING_BUDGET / STAGE_BYTES > 8 ? 8 : RING_BUDGET / STAGE_BYTES;
    static constexpr int TMEM_COLS = MX ? 512 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256 + (STAGED ? STG_BYTES : 0);
    static_assert(BN % 128 == 0 && BN <= 256, "BN must be 128 or 256 for CTA pairs");
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget exceeded");
};

// EB = bytes per input element: 2 = bf16 (kind::f16), 1 = fp8 e4m3 (kind::f8f6f4, K-major operands
// only). A K-block is always one 128-byte swizzle row, i.e. 64 bf16 or 128 fp8 elements, and an MMA
// always consumes 32 bytes of K, so the shared-memory descriptors and the pipeline are identical;
// only the instruction kind, the instruction descriptor and the K coordinates differ.
template <typename Policy>
__device__ __forceinline__ bool own_tile(const RankParams& p, const TileCoord& tc) {
    if constexpr (Policy::kOwnTileDirect) return Policy::is_own(p, tc);
    else return false;
}

// MX = block-scaled e4m3 (OCP MXFP8, one UE8M0 scale per 1 x 32 block; tcgen05.mma kind::mxf8f6f4.block_scale):
// EB = 1 plus two scale chunks per stage, copied shared -> tensor memory (tcgen05.cp) by the MMA thread right
// before the four MMAs of a K-block. Everything a Policy does -- gathers, scatters, epilogues -- is unchanged.
__device__ __forceinline__ void utccp_32x128b_2sm(uint32_t taddr, uint64_t sdesc) {
    asm volatile("tcgen05.cp.cta_group::2.32x128b.warpx4 [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_mx_ss_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t tmem_sfa, uint32_t tmem_sfb, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
        : "memory");
}

template <int BN, bool A_MN, bool B_MN, typename OutT, typename Policy, int EB = 2, bool MX = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_kernel_2cta(const __grid_constant__ Launch lp) {
    static_assert(EB == 2 || (EB == 1 && !A_MN && !B_MN), "fp8 operands must be K-major");
    static_assert(!MX || EB == 1, "block scaling is an fp8 mode");
    constexpr int KSCALE = 2 / EB;   // policies count K in bf16-sized elements (BK = 64 per block)
    using C = Cfg2<BN, Policy::kNeedsStaging, MX>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + C::STAGES * C::STAGE_BYTES);
    uint64_t* full = bars;
    uint64_t* empty = bars + C::STAGES;
    uint64_t* tmem_full = empty + C::STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    uint8_t* stg_base = reinterpret_cast<uint8_t*>(bars) + 256;
    auto smem_a = [&](int s) { return base + s * C::STAGE_BYTES; };
    auto smem_b = [&](int s) { return base + s * C::STAGE_BYTES + C::A_BYTES; };
    auto smem_sfa = [&](int s) { return base + s * C::STAGE_BYTES + C::A_BYTES + C::B_BYTES; };
    auto smem_sfb = [&](int s) { return base + s * C::STAGE_BYTES + C::A_BYTES + C::B_BYTES + C::SF_BYTES; };

    // clusters are consecutive block pairs; emulated ranks interleave at pair granularity
    const int nvr = lp.nvr;
    const int pair_global = blockIdx.x >> 1;
    const int vr = pair_global % nvr;
    const uint32_t cta_rank = ptx::cluster_ctarank();
    int cta = (pair_global / nvr) * 2 + (int)cta_rank;
    const int nctas_all = gridDim.x / nvr;
    const RankParams& p = lp.r[vr];
    const bool leader = cta_rank == 0;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    // device trace (RankParams::trace): each role's elected thread accumulates cycle counts in registers
    // and writes its slots once, at the end; with tracing off this costs a few predicated instructions
    unsigned long long* const trace = p.trace ? p.trace + (size_t)blockIdx.x * TRACE_WORDS : nullptr;
    if (trace && threadIdx.x == 0) trace[TR_T0_NS] = ptx::globaltimer_ns();

    if constexpr (Policy::kCommCtas) {
        const int ncomm = p.num_comm_ctas;  // even: whole clusters
        if (cta < ncomm) {
            Policy::comm_cta(p, cta, ncomm);
            return;
        }
        cta -= ncomm;
    }
    const int nctas = nctas_all - (Policy::kCommCtas ? p.num_comm_ctas : 0);
    const int pair = cta >> 1, npairs = nctas >> 1;

    const int m_blocks = (p.M + BM - 1) / BM;
    const int m_pairs = (m_blocks + 1) / 2;
    const int n_blocks = (p.N + BN - 1) / BN;
    const int k_blocks = (p.K + BK * KSCALE - 1) / (BK * KSCALE);
    const int ntiles = Policy::num_tiles(p, m_pairs, n_blocks);

    auto my_tile = [&](int t) {
        TileCoord tc = Policy::tile(p, t, m_pairs, n_blocks);
        tc.m_blk = tc.m_blk * 2 + (int)cta_rank;
        return tc;
    };

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&p.tmap_a);
        ptx::prefetch_tmap(&p.tmap_a2);
        ptx::prefetch_tmap(&p.tmap_b);
        if constexpr (MX) {
            ptx::prefetch_tmap(&p.tmap_sfa);
            ptx::prefetch_tmap(&p.tmap_sfa2);
            ptx::prefetch_tmap(&p.tmap_sfb);
        }
#pragma unroll
        for (int s = 0; s < C::STAGES; ++s) {
            ptx::mbar_init(&full[s], 2);
            ptx::mbar_init(&empty[s], 1);
        }
        ptx::mbar_init(&tmem_full[0], 1);
        ptx::mbar_init(&tmem_full[1], 1);
        ptx::mbar_init(&tmem_empty[0], 8);
        ptx::mbar_init(&tmem_empty[1], 8);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc_2sm<C::TMEM_COLS>(tmem_slot);
    ptx::tc_fence_before();
    ptx::cluster_sync();  // both CTAs' barriers exist before anyone signals across the pair
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            long long tr_flag = 0, tr_empty = 0, tr_max = 0, tr_slow = 0;
            unsigned long long tr_first = 0;
            const int nch = Policy::num_kchunks(p);
            const int per = k_blocks / nch;                       // K-blocks per chunk
            for (int t = pair; t < ntiles; t += npairs) {
                const TileCoord tc = my_tile(t);
                const bool valid = tc.m_blk < m_blocks;
                for (int ci = 0; ci < nch; ++ci) {
                    int k_base, a_k_base, a_m0;
                    const CUtensorMap* ta;
                    {
                        // a phantom m-block of an odd tail has rows out of range: TMA zero-fills them and there is
                        // nothing to wait for -- but the B half it loads feeds the partner's real rows, so K must
                        // advance in the policy's order
                        const long long c0 = trace ? clock64() : 0;
                        Policy::kchunk(p, tc, ci, valid, k_base, ta, a_k_base, a_m0);
                        if (trace) {
                            const long long d = clock64() - c0;
                            tr_flag += d;
                            if (d > tr_max) tr_max = d;
                            if (d > 2000) {
                                if (tr_slow == 0) tr_first = ptx::globaltimer_ns();
                                ++tr_slow;
                            }
                        }
                    }
                    for (int w = 0; w < per; ++w) {
                        const int k0 = (k_base + w * BK) * KSCALE;
                        const int a_k0 = (a_k_base + w * BK) * KSCALE;
                        {
                            const long long c0 = trace ? clock64() : 0;
                            ptx::mbar_wait(&empty[stage], phase ^ 1, 100 + stage);
                            if (trace) tr_empty += clock64() - c0;
                        }
                        if (leader) ptx::mbar_arrive_expect_tx(&full[stage], 2 * C::STAGE_BYTES);
                        else ptx::mbar_arrive_cluster(&full[stage], 0);
                        if constexpr (!A_MN) {
                            ptx::tma_load_2d_2sm(smem_a(stage), ta, &full[stage], a_k0, a_m0);
                        } else {
#pragma unroll
                            for (int c = 0; c < BM / 64; ++c)
                                ptx::tma_load_2d_2sm(smem_a(stage) + c * (BK * 128), ta, &full[stage], a_m0 + c * 64, a_k0);
                        }
                        const int n0 = tc.n_blk * BN + (int)cta_rank * C::B_ROWS;
                        if constexpr (!B_MN) {
                            ptx::tma_load_2d_2sm(smem_b(stage), &p.tmap_b, &full[stage], k0, n0);
                        } else {
#pragma unroll
                            for (int c = 0; c < C::B_ROWS / 64; ++c)
                                ptx::tma_load_2d_2sm(smem_b(stage) + c * (BK * 128), &p.tmap_b, &full[stage], n0 + c * 64, k0);
                        }
                        if constexpr (MX) {
                            // scale chunks: row `chunk` of a [chunks][128 x uint32] matrix, K-fastest chunk order. A's
                            // come from the array that belongs to the source of its rows (gathered staging or this
                            // rank's own shard), B's from the weight's. A phantom m-block reads the row past the end
                            // (TMA zero-fills it).
                            const bool local = ta == &p.tmap_a2;
                            const int sfa_row = (a_m0 / BM) * (local ? p.sf_a2_kblocks : k_blocks) + a_k_base / BK + w;
                            ptx::tma_load_2d_2sm(smem_sfa(stage), local ? &p.tmap_sfa2 : &p.tmap_sfa, &full[stage], 0, sfa_row);
                            ptx::tma_load_2d_2sm(smem_sfb(stage), &p.tmap_sfb, &full[stage], 0, tc.n_blk * k_blocks + k_base / BK + w);
                        }
                        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                    }
                }
            }
            if (trace) {
                trace[TR_PROD_FLAG] = (unsigned long long)tr_flag;
                trace[TR_PROD_EMPTY] = (unsigned long long)tr_empty;
                trace[TR_PROD_SLOW] = (unsigned long long)tr_slow;
                trace[TR_PROD_MAX] = (unsigned long long)tr_max;
                trace[TR_PROD_FIRST_NS] = tr_first;
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (leader && lane == 0) {
            constexpr uint32_t idesc = EB == 2 ? ptx::umma_idesc(1, 1, A_MN, B_MN, 2 * BM, BN)     // bf16 x bf16
                                               : ptx::umma_idesc(0, 0, false, false, 2 * BM, BN);  // e4m3 x e4m3
            int stage = 0;
            uint32_t phase = 0, kb = 0;
            int local = 0;
            long long tr_full = 0, tr_acc = 0;
            for (int t = pair; t < ntiles; t += npairs, ++local) {
                const int acc = local & 1;
                const uint32_t acc_phase = (local >> 1) & 1;
                {
                    const long long c0 = trace ? clock64() : 0;
                    ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1, 200 + acc);
                    if (trace) tr_acc += clock64() - c0;
                }
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int it = 0; it < k_blocks; ++it, ++kb) {
                    {
                        const long long c0 = trace ? clock64() : 0;
                        ptx::mbar_wait(&full[stage], phase, 300 + stage);
                        if (trace) tr_full += clock64() - c0;
                    }
                    ptx::tc_fence_after();
                    const uint32_t a_addr = ptx::smem_u32(smem_a(stage));
                    const uint32_t b_addr = ptx::smem_u32(smem_b(stage));
                    if constexpr (MX) {
                        // scale columns are double-buffered by K-block parity; tcgen05.cp and tcgen05.mma execute in
                        // issue order, so the copies need no barrier of their own
                        const uint32_t sf = tmem_base + C::SF_COL0 + (kb & 1) * 8;     // 4 columns SFA, then 4 columns SFB
                        utccp_32x128b_2sm(sf, mx::smem_desc_noswizzle(ptx::smem_u32(smem_sfa(stage)), mx::SF_CHUNK_LBO, mx::SF_CHUNK_SBO));
                        utccp_32x128b_2sm(sf + 4, mx::smem_desc_noswizzle(ptx::smem_u32(smem_sfb(stage)), mx::SF_CHUNK_LBO, mx::SF_CHUNK_SBO));
#pragma unroll
                        for (uint32_t ks = 0; ks < 4; ++ks)
                            umma_mx_ss_2sm(tmem_d, operand_desc<false>(a_addr, ks), operand_desc<false>(b_addr, ks),
                                           mx::idesc_mx(2 * BM, BN, ks), sf | (ks << 30), (sf + 4) | (ks << 30),
                                           (it | (int)ks) != 0 ? 1u : 0u);
                        ptx::umma_commit_2sm(&empty[stage], 0b11);
                        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
#pragma unroll
                    for (int ks = 0; ks < BK / UMMA_K; ++ks) {
                        if constexpr (EB == 2)
                            ptx::umma_f16_ss_2sm(tmem_d, operand_desc<A_MN>(a_addr, ks), operand_desc<B_MN>(b_addr, ks),
                                                 idesc, (it | ks) != 0 ? 1u : 0u);
                        else
                            ptx::umma_f8_ss_2sm(tmem_d, operand_desc<A_MN>(a_addr, ks), operand_desc<B_MN>(b_addr, ks),
                                                idesc, (it | ks) != 0 ? 1u : 0u);
                    }
                    ptx::umma_commit_2sm(&empty[stage], 0b11);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit_2sm(&tmem_full[acc], 0b11);
            }
            if (trace) {
                trace[TR_MMA_FULL] = (unsigned long long)tr_full;
                trace[TR_MMA_ACC] = (unsigned long long)tr_acc;
            }
        }
        __syncwarp();
    } else {
        // ===================== epilogue warps (both CTAs, own 128 rows) =====================
        const int q = warp & 3;
        const int epi_tid = (warp - EPI_WARP0) * 32 + lane;
        int local = 0;
        typename Policy::State st;
        volatile uint32_t* bcast = tmem_slot + 2;          // two words: ping-pong decision broadcast
        Policy::init_state(st, p, cta, nctas, epi_tid,
                           Policy::kNeedsStaging ? ptx::smem_u32(stg_base) + epi_tid * STG_ROW_STRIDE : 0u,
                           reinterpret_cast<uint64_t*>(tmem_slot + 8) + (warp - EPI_WARP0));
        Policy::prologue(p, cta, nctas, epi_tid);
        if constexpr (Policy::kBackground) {
            // diagnostic switch (aux1 bit 2): do all pushes up front instead of in the gaps
            if (Policy::kPushFirst && (p.aux1 & 4))
                while (Policy::template background<OutT>(p, st, cta, nctas, epi_tid, bcast + 2)) {}
        }
        uint32_t poll = 0;
        long long tr_bg = 0, tr_wait = 0, tr_slices = 0;
        const bool tr_on = trace != nullptr && epi_tid == 0;
        for (int t = pair; t < ntiles; t += npairs, ++local) {
            const TileCoord tc = my_tile(t);
            const bool valid = tc.m_blk < m_blocks;
            const int acc = local & 1;
            const uint32_t acc_phase = (local >> 1) & 1;
            const int row = tc.m_blk * BM + q * 32 + lane;
            // second operand of a fused elementwise epilogue: the first chunk's 64 bytes are requested
            // before the wait for the accumulator, every further chunk one chunk ahead of its use
            uint4 epi_nxt[4];
            bool epi_in = false;
            if constexpr (Policy::kEpiOps && !Policy::kStagedStore) {
                epi_in = valid && Policy::epi_has_input(p);
                if (epi_in) Policy::epi_load(p, row, tc.n_blk * BN, epi_nxt);
            }
            const long long c_bg = tr_on ? clock64() : 0;
            if constexpr (Policy::kBackground) {
                // While the MMAs of this tile are still running, spend the time on the policy's
                // background work, one bounded slice at a time; the decision is taken by one
                // thread and shared, because slices contain barriers of the 128 epilogue threads.
                for (;;) {
                    volatile uint32_t* w = bcast + (poll++ & 1);
                    if (epi_tid == 0) *w = ptx::mbar_test_wait(&tmem_full[acc], acc_phase) ? 1u : 0u;
                    epi_bar_sync();
                    if (*w != 0) break;
                    if (!Policy::template background<OutT>(p, st, cta, nctas, epi_tid, bcast + 2)) break;
                    ++tr_slices;
                }
            }
            const long long c_w = tr_on ? clock64() : 0;
            ptx::mbar_wait(&tmem_full[acc], acc_phase, 400 + acc);
            if (tr_on) {
                const long long c1 = clock64();
                tr_bg += c_w - c_bg;
                tr_wait += c1 - c_w;
            }
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * BN;
            if constexpr (!Policy::kStagedStore) {
#pragma unroll 1
                for (int c = 0; c < BN; c += 32) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32b_x32(taddr + c, v);
                    uint4 epi_cur[4];
                    if constexpr (Policy::kEpiOps) {
                        if (epi_in) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) epi_cur[j] = epi_nxt[j];
                            if (c + 32 < BN) Policy::epi_load(p, row, tc.n_blk * BN + c + 32, epi_nxt);
                        }
                    }
                    ptx::tmem_ld_wait();
                    if (c + 32 >= BN) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if (leader) ptx::mbar_arrive(&tmem_empty[acc]);
                            else ptx::mbar_arrive_cluster(&tmem_empty[acc], 0);
                        }
                    }
                    if constexpr (Policy::kEpiOps && sizeof(OutT) == 2) {
                        if (valid && p.epi_op != EPI_NONE) {
                            Policy::template store_epi<OutT>(p, tc, row, tc.n_blk * BN + c, v, epi_cur);
                            continue;
                        }
                    }
                    if (valid) Policy::template store<OutT>(p, tc, row, tc.n_blk * BN + c, v);
                }
            } else if (Policy::kOwnTileDirect && valid && own_tile<Policy>(p, tc)) {
                // a tile this rank owns (reduce-scatter): add the peers' partials and store the final values
                if constexpr (Policy::kOwnTileDirect) {
                    Policy::own_wait(p, tc, epi_tid);
                    typename Policy::template OwnPf<OutT> pf_nxt;
                    Policy::template own_prefetch<OutT>(p, row, tc.n_blk * BN, pf_nxt);
#pragma unroll 1
                    for (int c = 0; c < BN; c += 32) {
                        uint32_t v[32];
                        ptx::tmem_ld_32x32b_x32(taddr + c, v);
                        const typename Policy::template OwnPf<OutT> pf_cur = pf_nxt;
                        if (c + 32 < BN) Policy::template own_prefetch<OutT>(p, row, tc.n_blk * BN + c + 32, pf_nxt);
                        ptx::tmem_ld_wait();
                        if (c + 32 >= BN) {
                            ptx::tc_fence_before();
                            __syncwarp();
                            if (lane == 0) {
                                if (leader) ptx::mbar_arrive(&tmem_empty[acc]);
                                else ptx::mbar_arrive_cluster(&tmem_empty[acc], 0);
                            }
                        }
                        Policy::template own_store<OutT>(p, tc, row, tc.n_blk * BN + c, v, pf_cur);
                    }
                }
            } else {
                constexpr int CW = STG_CHUNK_BYTES / (int)sizeof(OutT);   // columns per staged chunk
                const uint32_t my_row = ptx::smem_u32(stg_base) + ((warp - EPI_WARP0) * 32 + lane) * STG_ROW_STRIDE;
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += CW) {
                    ptx::tma_store_wait_read<0>();   // this lane's previous copies have left its row buffer
#pragma unroll 1
                    for (int c = c0; c < c0 + CW; c += 32) {
                        uint32_t v[32];
                        ptx::tmem_ld_32x32b_x32(taddr + c, v);
                        ptx::tmem_ld_wait();
                        if (c + 32 >= BN) {
                            ptx::tc_fence_before();
                            __syncwarp();
                            if (lane == 0) {
                                if (leader) ptx::mbar_arrive(&tmem_empty[acc]);
                                else ptx::mbar_arrive_cluster(&tmem_empty[acc], 0);
                            }
                        }
                        const uint32_t dst = my_row + (c - c0) * (int)sizeof(OutT);
                        if constexpr (sizeof(OutT) == 2) {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                ptx::st_shared_v4(dst + 16 * j, make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                                                                           pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7])));
                        } else {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                ptx::st_shared_v4(dst + 16 * j, make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
                        }
                    }
                    ptx::fence_proxy_async_smem();   // own generic writes -> own bulk copy (async proxy)
                    const int col0 = tc.n_blk * BN + c0;
                    if (valid && row < p.M && col0 < p.N) {
                        const int ncols = min(CW, p.N - col0);
                        Policy::template bulk_row<OutT>(p, tc, row, col0, my_row, ncols * (int)sizeof(OutT));
                    }
                    ptx::tma_store_commit();
                }
                ptx::tma_store_wait<0>();            // this lane's rows have landed
                ptx::fence_proxy_async_all();        // ... and are ordered before the flag traffic below
            }
            if (valid) Policy::tile_done(p, tc, epi_tid);
        }
        const long long c_tail = tr_on ? clock64() : 0;
        Policy::template tail<OutT>(p, st, cta, nctas, epi_tid, bcast + 2);
        if (tr_on) {
            trace[TR_EPI_BG] = (unsigned long long)tr_bg;
            trace[TR_EPI_ACC] = (unsigned long long)tr_wait;
            trace[TR_EPI_TAIL] = (unsigned long long)(clock64() - c_tail);
            trace[TR_T1_NS] = ptx::globaltimer_ns();
            trace[TR_TILES] = (unsigned long long)local;
            trace[TR_EPI_SLICES] = (unsigned long long)tr_slices;
        }
    }

    // ---- teardown: nobody leaves (or frees TMEM) while the partner may still read its smem ----
    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_2sm<C::TMEM_COLS>(tmem_base);
    }
}

}  // namespace gemm
