// Persistent, warp-specialised tcgen05 GEMM mainloop for sm_100a, shared by the plain GEMM and
// the fused GEMM+collective kernels.
//
//   warp 0 lane 0 : TMA producer  (global -> 128B-swizzled smem ring, mbarrier tx-count)
//   warp 1 lane 0 : MMA issuer    (tcgen05.mma 128 x BN x 16, fp32 accumulators in TMEM)
//   warps 2..5    : epilogue      (tcgen05.ld -> convert -> policy-defined stores)
//
// Two accumulator stages (2 x BN TMEM columns) let the epilogue of tile i overlap the mainloop
// of tile i+1. "Policies" customise, at compile time, (a) the order in which tiles and K-blocks
// are visited and what the producer must wait for before a load, and (b) where the epilogue
// puts a finished tile. That is the seam where the collectives of the reference
// (func_impl.py:89 all-gather, :107 all-gather, :184-186 all-to-all + sum) get fused into the
// FC GEMMs the reference leaves out (README.md:173-175).
#pragma once
#include "ptx.cuh"
#include "../include/ccmpi_b200.h"

namespace gemm {

constexpr int BM = 128;      // UMMA M (one CTA, all 128 TMEM lanes)
constexpr int BK = 64;       // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;   // fixed for 16-bit inputs
constexpr int NUM_THREADS = 192;
constexpr int EPI_WARP0 = 2;  // first epilogue warp
constexpr int EPI_THREADS = 128;
// staging area of the CTA-pair kernel's epilogue warps: one 256-byte row (+16 bytes of padding) per lane,
// i.e. 8.5 KB contiguous per warp
constexpr int STG_CHUNK_BYTES = 256;
constexpr int STG_ROW_STRIDE = STG_CHUNK_BYTES + 16;
constexpr int STG_BYTES = 4 * 32 * STG_ROW_STRIDE;       // 4 epilogue warps x 32 rows

template <int BN>
struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;   // 16 KiB
    static constexpr int B_BYTES = BN * BK * 2;   // 32 KiB at BN=256
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (BN >= 256) ? 4 : ((BN >= 128) ? 6 : 8);
    static constexpr int TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
    static_assert(BN % 64 == 0 && BN <= 256, "BN must be 64, 128, 192 or 256");
};

struct TileCoord {
    int m_blk, n_blk;
};

// One rank's view of a GEMM problem. Lives in kernel parameter space (grid constant) because
// TMA descriptors must be read through the param/const/global window.
struct alignas(64) RankParams {
    CUtensorMap tmap_a;   // main A operand
    CUtensorMap tmap_a2;  // alternate A source (fused all-gather: the rank-local shard)
    CUtensorMap tmap_b;
    void* C;              // output (policy-defined meaning)
    long long ldc;
    int M, N, K;
    int accumulate;       // C += A*B (fp32 output only)
    // ---- fused-collective fields (unused by the plain policy) ----
    ccmpi_dev_comm comm;
    unsigned long long off0, off1, off2;  // heap offsets, policy-defined
    int chunk;            // per-rank chunk extent along the gathered / scattered axis
    int flag_base;        // first user flag word used by this launch
    unsigned int epoch;   // value flags must reach for this launch
    int num_comm_ctas;    // CTAs (per rank) that move data instead of computing tiles
    int aux0, aux1, aux2;
    // ---- elementwise work fused into the epilogue (bf16 outputs; see EpiOp) ----
    int epi_op;
    const void* epi_in;   // second epilogue operand  [M][epi_ld] bf16: u for EPI_MUL_DGELU, the residual for EPI_ADD
    void* epi_out2;       // second output            [M][epi_ld] bf16: gelu(C) for EPI_GELU_DUAL
    long long epi_ld;
    // ---- block-scaled (MXFP8) operands: scale chunks of A (gathered staging / this rank's own shard) and of B ----
    CUtensorMap tmap_sfa, tmap_sfa2, tmap_sfb;
    int sf_a2_kblocks;    // K-blocks per 128-row chunk row of the tmap_sfa2 array
    int sf_run_bytes;     // all-gather: bytes of this rank's scale run per m-block (0: no scales to push)
    unsigned long long off_sf_local, off_sf_full;   // symmetric offsets of the local / gathered scale arrays
    // ---- per-CTA device trace (nullptr: off): TRACE_WORDS x uint64 per CTA, see TraceSlot ----
    unsigned long long* trace;
};

// Elementwise work a GEMM can do on its accumulator tile before it leaves the SM, so that the
// activation / residual passes between the FC GEMMs cost no extra trip through HBM.
enum EpiOp : int {
    EPI_NONE = 0,
    EPI_GELU_DUAL = 1,   // C = acc (the pre-activation, kept for backward) and epi_out2 = gelu(acc)
    EPI_MUL_DGELU = 2,   // C = acc * gelu'(epi_in)      (backward through the activation)
    EPI_ADD = 3,         // C = acc + epi_in             (residual connection / gradient accumulation)
};

// What each CTA records when RankParams::trace is set (SM cycles unless noted). Written once, at the end,
// by the elected thread of each role; read on the host by utils/ktrace.py.
enum TraceSlot : int {
    TR_T0_NS = 0,        // %globaltimer at kernel entry
    TR_PROD_FLAG = 1,    // TMA producer: cycles inside Policy::kblock (waits for gathered chunks)
    TR_PROD_EMPTY = 2,   // TMA producer: cycles waiting for a free pipeline stage (MMA-bound: good)
    TR_MMA_FULL = 3,     // MMA issuer: cycles waiting for operands (starved)
    TR_MMA_ACC = 4,      // MMA issuer: cycles waiting for a free accumulator (epilogue-bound)
    TR_EPI_BG = 5,       // epilogue: cycles in background work (pushes, reductions)
    TR_EPI_ACC = 6,      // epilogue: cycles waiting for a finished accumulator
    TR_EPI_TAIL = 7,     // epilogue: cycles in the policy's tail (left-over pushes / reductions / barriers)
    TR_T1_NS = 8,        // %globaltimer when the epilogue warps are done
    TR_TILES = 9,        // tiles this CTA computed
    TR_PROD_SLOW = 10,   // TMA producer: kblock calls that took more than ~1 us
    TR_PROD_MAX = 11,    // TMA producer: longest single kblock call
    TR_PROD_FIRST_NS = 12, // TMA producer: ns from kernel entry until its first slow kblock call returned
    TR_EPI_SLICES = 13,  // epilogue: background slices executed
    TRACE_WORDS = 16
};

struct Launch {
    int nvr;  // emulated ranks in this launch (1 in production)
    int pad_[15];
    RankParams r[CCMPI_MAX_VRANKS];
};

// ------------------------------------------------------------------------------------------------
// Shared-memory carve-up
// ------------------------------------------------------------------------------------------------
template <int BN>
struct Smem {
    uint8_t* base;         // 1024-byte aligned start of the stage ring
    uint64_t* full;        // [STAGES]  TMA -> MMA
    uint64_t* empty;       // [STAGES]  MMA -> TMA
    uint64_t* tmem_full;   // [2]       MMA -> epilogue
    uint64_t* tmem_empty;  // [2]       epilogue -> MMA
    uint32_t* tmem_base;   // TMEM allocation result

    __device__ explicit Smem(uint8_t* raw) {
        base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
        uint64_t* bars = reinterpret_cast<uint64_t*>(base + Cfg<BN>::STAGES * Cfg<BN>::STAGE_BYTES);
        full = bars;
        empty = bars + Cfg<BN>::STAGES;
        tmem_full = empty + Cfg<BN>::STAGES;
        tmem_empty = tmem_full + 2;
        tmem_base = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    }
    __device__ __forceinline__ uint8_t* a(int s) const { return base + s * Cfg<BN>::STAGE_BYTES; }
    __device__ __forceinline__ uint8_t* b(int s) const { return base + s * Cfg<BN>::STAGE_BYTES + Cfg<BN>::A_BYTES; }
};

// ------------------------------------------------------------------------------------------------
// Operand staging: how a (rows x 64) K-block of A or B gets from global to smem and how the
// UMMA descriptor walks it. K-major: one TMA box [64 k][rows]. MN-major (operand stored
// [K][MN]): rows/64 boxes of [64 mn][64 k], each an 8 KiB slab.
// ------------------------------------------------------------------------------------------------
template <bool MN_MAJOR, int ROWS>
__device__ __forceinline__ void load_operand(uint8_t* smem, const CUtensorMap* tmap, uint64_t* bar,
                                             int k0, int mn0) {
    if constexpr (!MN_MAJOR) {
        ptx::tma_load_2d(smem, tmap, bar, k0, mn0);
    } else {
#pragma unroll
        for (int c = 0; c < ROWS / 64; ++c)
            ptx::tma_load_2d(smem + c * (BK * 128), tmap, bar, mn0 + c * 64, k0);
    }
}

template <bool MN_MAJOR>
__device__ __forceinline__ uint64_t operand_desc(uint32_t smem_addr, int kstep) {
    if constexpr (!MN_MAJOR) {
        // rows are 128 B apart, 8-row swizzle atoms are 1024 B apart; K advances inside the row
        return ptx::umma_smem_desc_sw128(smem_addr + kstep * (UMMA_K * 2), 0, 1024);
    } else {
        // 64-wide MN slabs are BK*128 B apart (LBO); 8-k-row atoms are 1024 B apart (SBO)
        return ptx::umma_smem_desc_sw128(smem_addr + kstep * (UMMA_K * 128), BK * 128, 1024);
    }
}

// Grouped rasterisation: walk 8 m-blocks under each n-block before moving on, so the CTAs that
// run concurrently share A row-panels and B column-panels in L2.
__device__ __forceinline__ TileCoord raster(int tile, int m_blocks, int n_blocks) {
    constexpr int GROUP_M = 8;
    const int per_group = GROUP_M * n_blocks;
    const int g = tile / per_group;
    const int first_m = g * GROUP_M;
    const int gsize = min(m_blocks - first_m, GROUP_M);
    const int in_g = tile - g * per_group;
    return TileCoord{first_m + in_g % gsize, in_g / gsize};
}

}  // namespace gemm
