// Block-scaled fp8 GEMM (OCP MXFP8: e4m3 values, one UE8M0 scale per 1 x 32 block) on CTA pairs:
//     C[M][N] = (A * 2^sfa) (B * 2^sfb)^T        A: [M][K] e4m3, B: [N][K] e4m3, both K-major
// with tcgen05.mma.cta_group::2.kind::mxf8f6f4.block_scale -- the tensor cores apply the scales, which
// sit in tensor memory next to the accumulators (DESIGN.md section 6 has the facts this file relies on).
// This is SURVEY K14's GEMM; the reference has no GEMM at all.
//
// Structure = the bf16 CTA-pair kernel (gemm_kernel_2cta.cuh) with two additions:
//   * every pipeline stage also carries one 512-byte scale chunk for this CTA's 128 rows of A and one for
//     the tile's 128 columns of B (TMA, no swizzle, from the chunked layout ops/gemm_fp8.py produces);
//   * before the four MMAs of a K-block the issuing thread copies both chunks shared -> tensor memory with
//     tcgen05.cp (32 lanes x 128 bit, broadcast to the four lane quadrants); tcgen05.cp and tcgen05.mma
//     execute in issue order, and the scale columns are double-buffered by K-block parity.
// Tile: 256 x 128 per pair (two 128-column accumulators + 16 scale columns of tensor memory).
//
// Run on a B200 in round 2 (tests/test_gpu_fp8.py); the MX mode of the generic CTA-pair kernel agrees with it
// bit for bit (tests/test_gpu_mx_fused.py). Checked without a GPU as well: the instruction descriptor, the
// shared-memory descriptor of a scale chunk (LBO 16 B, SBO 128 B, no swizzle) and the scale layouts in shared
// and tensor memory (32 lanes x 4 columns per chunk, scale id = byte of the column, replicated to the four lane
// quadrants) are bit-for-bit what the CUTLASS 4.5 headers on the image compute for this tile (evaluated in a
// host program, scripts/cutlass_crosscheck.sh).
#include "gemm_host.h"
#include "mx_desc.cuh"

namespace mx {

using namespace gemm;

constexpr int BN = 128;
constexpr int A_BYTES = BM * 128;            // 128 rows x 128 fp8
constexpr int B_ROWS = BN / 2;               // this CTA's half of the B tile
constexpr int B_BYTES = B_ROWS * 128;
constexpr int SF_BYTES = 512;                // one chunk: 128 rows (or columns) x 4 scales
constexpr int STAGE_BYTES = A_BYTES + B_BYTES + 2 * SF_BYTES;    // 25600 = 25 x 1024
constexpr int STAGES = 7;
constexpr int TMEM_COLS = 512;               // 2 x 128 accumulator columns + scales, rounded to a power of two
constexpr int SF_COL0 = 2 * BN;              // first scale column
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
static_assert(STAGE_BYTES % 1024 == 0, "operand tiles must stay 1024-byte aligned");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget exceeded");

struct Params {
    CUtensorMap tmap_a, tmap_b, tmap_sfa, tmap_sfb;
    void* C;
    long long ldc;
    int M, N, K;
    int out_bf16;
};

// 32 rows x 16 bytes of shared memory -> 4 columns of tensor memory, replicated to the four lane quadrants
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

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1) gemm_mxfp8_kernel(const __grid_constant__ Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(base + STAGES * STAGE_BYTES);
    uint64_t* full = bars;
    uint64_t* empty = bars + STAGES;
    uint64_t* tmem_full = empty + STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    auto smem_a = [&](int s) { return base + s * STAGE_BYTES; };
    auto smem_b = [&](int s) { return base + s * STAGE_BYTES + A_BYTES; };
    auto smem_sfa = [&](int s) { return base + s * STAGE_BYTES + A_BYTES + B_BYTES; };
    auto smem_sfb = [&](int s) { return base + s * STAGE_BYTES + A_BYTES + B_BYTES + SF_BYTES; };

    const uint32_t cta_rank = ptx::cluster_ctarank();
    const bool leader = cta_rank == 0;
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const int m_blocks = (p.M + BM - 1) / BM;
    const int m_pairs = (m_blocks + 1) / 2;
    const int n_blocks = (p.N + BN - 1) / BN;
    const int k_blocks = (p.K + 127) / 128;           // 128 fp8 elements = one swizzle row = 4 scales
    const int ntiles = m_pairs * n_blocks;

    auto my_tile = [&](int t) {
        TileCoord tc = raster(t, m_pairs, n_blocks);
        tc.m_blk = tc.m_blk * 2 + (int)cta_rank;
        return tc;
    };

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&p.tmap_a);
        ptx::prefetch_tmap(&p.tmap_b);
        ptx::prefetch_tmap(&p.tmap_sfa);
        ptx::prefetch_tmap(&p.tmap_sfb);
#pragma unroll
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full[s], 2);
            ptx::mbar_init(&empty[s], 1);
        }
        ptx::mbar_init(&tmem_full[0], 1);
        ptx::mbar_init(&tmem_full[1], 1);
        ptx::mbar_init(&tmem_empty[0], 8);
        ptx::mbar_init(&tmem_empty[1], 8);
        ptx::fence_mbar_init();
    }
    if (warp == 1) ptx::tmem_alloc_2sm<TMEM_COLS>(tmem_slot);
    ptx::tc_fence_before();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs): A rows, B half, both scale chunks =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = pair; t < ntiles; t += npairs) {
                const TileCoord tc = my_tile(t);
                for (int it = 0; it < k_blocks; ++it) {
                    ptx::mbar_wait(&empty[stage], phase ^ 1, 100 + stage);
                    if (leader) ptx::mbar_arrive_expect_tx(&full[stage], 2 * STAGE_BYTES);
                    else ptx::mbar_arrive_cluster(&full[stage], 0);
                    ptx::tma_load_2d_2sm(smem_a(stage), &p.tmap_a, &full[stage], it * 128, tc.m_blk * BM);
                    ptx::tma_load_2d_2sm(smem_b(stage), &p.tmap_b, &full[stage], it * 128,
                                         tc.n_blk * BN + (int)cta_rank * B_ROWS);
                    // scale chunks: row `chunk` of a [chunks][128 x uint32] matrix, K-fastest chunk order.
                    // A phantom m-block of an odd tail reads the (neutral, padded) chunk row past the end.
                    ptx::tma_load_2d_2sm(smem_sfa(stage), &p.tmap_sfa, &full[stage], 0, tc.m_blk * k_blocks + it);
                    ptx::tma_load_2d_2sm(smem_sfb(stage), &p.tmap_sfb, &full[stage], 0, tc.n_blk * k_blocks + it);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (leader && lane == 0) {
            int stage = 0;
            uint32_t phase = 0, kb = 0;
            int local = 0;
            for (int t = pair; t < ntiles; t += npairs, ++local) {
                const int acc = local & 1;
                const uint32_t acc_phase = (local >> 1) & 1;
                ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1, 200 + acc);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int it = 0; it < k_blocks; ++it, ++kb) {
                    ptx::mbar_wait(&full[stage], phase, 300 + stage);
                    ptx::tc_fence_after();
                    const uint32_t sf = tmem_base + SF_COL0 + (kb & 1) * 8;     // 4 columns SFA, then 4 columns SFB
                    // LBO 16 B / SBO 128 B: the values CUTLASS's get_utccp_smem_desc_tensor produces for this chunk
                    utccp_32x128b_2sm(sf, smem_desc_noswizzle(ptx::smem_u32(smem_sfa(stage)), SF_CHUNK_LBO, SF_CHUNK_SBO));
                    utccp_32x128b_2sm(sf + 4, smem_desc_noswizzle(ptx::smem_u32(smem_sfb(stage)), SF_CHUNK_LBO, SF_CHUNK_SBO));
                    const uint32_t a_addr = ptx::smem_u32(smem_a(stage));
                    const uint32_t b_addr = ptx::smem_u32(smem_b(stage));
#pragma unroll
                    for (uint32_t ks = 0; ks < 4; ++ks) {
                        // scale id = byte of the 32-bit tensor-memory word = top two bits of the scale address
                        umma_mx_ss_2sm(tmem_d, operand_desc<false>(a_addr, ks), operand_desc<false>(b_addr, ks),
                                       idesc_mx(2 * BM, BN, ks), sf | (ks << 30), (sf + 4) | (ks << 30),
                                       (it | (int)ks) != 0 ? 1u : 0u);
                    }
                    ptx::umma_commit_2sm(&empty[stage], 0b11);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit_2sm(&tmem_full[acc], 0b11);
            }
        }
        __syncwarp();
    } else {
        // ===================== epilogue warps (both CTAs, own 128 rows) =====================
        const int q = warp & 3;
        int local = 0;
        for (int t = pair; t < ntiles; t += npairs, ++local) {
            const TileCoord tc = my_tile(t);
            const int acc = local & 1;
            const uint32_t acc_phase = (local >> 1) & 1;
            ptx::mbar_wait(&tmem_full[acc], acc_phase, 400 + acc);
            ptx::tc_fence_after();
            const int row = tc.m_blk * BM + q * 32 + lane;
            const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
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
                const int col0 = tc.n_blk * BN + c;
                const int ncols = min(32, p.N - col0);
                if (row < p.M && ncols > 0) {
                    if (p.out_bf16)
                        store_row32<__nv_bfloat16>(reinterpret_cast<__nv_bfloat16*>(p.C) + (long long)row * p.ldc + col0, v, ncols, false);
                    else
                        store_row32<float>(reinterpret_cast<float*>(p.C) + (long long)row * p.ldc + col0, v, ncols, false);
                }
            }
        }
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_2sm<TMEM_COLS>(tmem_base);
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// scale buffer as a [chunks][128] matrix of 32-bit words, one 512-byte chunk per TMA box, no swizzle
static int encode_sf_map(CUtensorMap* out, const void* base, long long chunks) {
    auto fn = ccmpi::drv<EncodeTiledFn>("cuTensorMapEncodeTiled");
    if (!fn) return CCMPI_ERR_CUDA;
    CCMPI_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "scale buffer %p is not 16-byte aligned", base);
    const cuuint64_t dims[2] = {128, (cuuint64_t)chunks};
    const cuuint64_t strides[1] = {512};
    const cuuint32_t box[2] = {128, 1};
    const cuuint32_t ones[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, ones,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        ccmpi::set_error("cuTensorMapEncodeTiled (scales) failed: CUresult %d", (int)r);
        return CCMPI_ERR_CUDA;
    }
    return CCMPI_OK;
}

}  // namespace mx

// sfa / sfb: scales in the chunked layout (512-byte chunks of 128 rows x 4 scales, K-fastest; see
// ops/gemm_fp8.py::swizzle_scales), padded to whole chunks: ceil(M/128) * K/128 and ceil(N/128) * K/128 chunks.
extern "C" int ccmpi_gemm_mxfp8(const void* A, const void* sfa, const void* B, const void* sfb, void* C, int M, int N,
                                int K, int lda, int ldb, int ldc, int out_dtype, int num_sms, void* stream) {
    using namespace mx;
    CCMPI_REQUIRE(A && B && C && sfa && sfb, "null operand");
    CCMPI_REQUIRE(M > 0 && N > 0 && K > 0 && M % 8 == 0 && N % 8 == 0 && K % 128 == 0,
                  "mxfp8 GEMM extents: M, N multiples of 8 and K a multiple of 128 (M=%d N=%d K=%d)", M, N, K);
    CCMPI_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && ldc % 8 == 0, "mxfp8 GEMM leading dimensions must keep rows 16-byte aligned");
    CCMPI_REQUIRE(out_dtype == CCMPI_BF16 || out_dtype == CCMPI_F32, "mxfp8 GEMM output must be bf16 or fp32");
    const int sms = (num_sms > 0 ? num_sms : ccmpi::device_sm_count()) & ~1;
    CCMPI_REQUIRE(sms >= 2, "need at least one SM pair");
    Params p;
    memset(&p, 0, sizeof(p));
    int rc = ccmpi::make_tmap_2d(&p.tmap_a, A, 1, M, K, lda, BM, 128);
    if (rc) return rc;
    rc = ccmpi::make_tmap_2d(&p.tmap_b, B, 1, N, K, ldb, B_ROWS, 128);
    if (rc) return rc;
    const long long kc = K / 128;
    const long long m_chunks = (M + 127) / 128, n_chunks = (N + 127) / 128;
    // the phantom m-block of an odd number of m-blocks reads one chunk row past the end: TMA zero-fills it
    rc = encode_sf_map(&p.tmap_sfa, sfa, m_chunks * kc);
    if (rc) return rc;
    rc = encode_sf_map(&p.tmap_sfb, sfb, n_chunks * kc);
    if (rc) return rc;
    p.C = C;
    p.ldc = ldc;
    p.M = M; p.N = N; p.K = K;
    p.out_bf16 = out_dtype == CCMPI_BF16;
    const int pair_tiles = (int)(((m_chunks + 1) / 2) * n_chunks);
    const int npairs = pair_tiles < sms / 2 ? pair_tiles : sms / 2;
    static bool configured[64] = {};   // the attribute is per device (and per instantiation)
    int dev_ = 0;
    CCMPI_CUDA_CHECK(cudaGetDevice(&dev_));
    if (!configured[dev_ & 63]) {
        CCMPI_CUDA_CHECK(cudaFuncSetAttribute(gemm_mxfp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        configured[dev_ & 63] = true;
    }
    gemm_mxfp8_kernel<<<2 * npairs, NUM_THREADS, SMEM_BYTES, (cudaStream_t)stream>>>(p);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}
