// ccmpi_gemm_rs_local_noalloc: the GEMM -> reduce-scatter kernel for a ONE-GPU communicator,
// with every byte of its scratch memory in a file-scope static device array.
//
// Who this is for. Some hosts of this library may not allocate device memory at the moment they
// need a matmul: an inference server that handed the whole GPU to its own arena at start-up and
// treats any later cudaMalloc as a bug; a plug-in running inside another framework's operator,
// where memory comes from that framework's planner and a driver allocation would bypass (and
// possibly deadlock with) it; a latency-critical loop that cannot afford cudaMalloc's implicit
// device-wide synchronisation. The regular entry point, ccmpi_fused_gemm_rs, needs a staging
// buffer ("slots", one partial tile set per source rank) and flag words that somebody has to
// allocate -- normally the symmetric heap. This variant needs nobody to: the caller passes only
// A, B and C.
//
// What it holds. `g_scratch` below: 64 KiB of flag words followed by CCMPI_NOALLOC_MAX_BYTES of
// staging. It is part of the library image, so EVERY GPU that loads libccmpi_b200.so carries it
// from load until the process exits, used or not.
//
// Rules that follow from sharing one array:
//  * All calls on a device use the same scratch and it keeps whatever the previous call left.
//    The flag words must start at zero, so each call zeroes them on the caller's stream before
//    the kernel; the staging area is fully overwritten before it is read.
//  * Two calls that overlap on the GPU (different streams, or different host threads) race on the
//    scratch and produce garbage. Serialise them.
//  * The first call on a device looks up the array's device address (cudaGetSymbolAddress). Make
//    that first call before any stream capture on that device.
//  * M * N * sizeof(output element) must not exceed CCMPI_NOALLOC_MAX_BYTES; larger problems are
//    refused with an error, never truncated.
// Nothing here calls cudaMalloc / cudaFree or any library routine that allocates device memory;
// TMA descriptors are encoded on the host stack and passed by value in the kernel parameters.
#include <mutex>

#include "fused_policies.cuh"
#include "host_common.h"

#define CCMPI_NOALLOC_PAD_BYTES (CCMPI_PAD_WORDS * 4)
#define CCMPI_NOALLOC_MAX_BYTES (16u << 20)

// In host code this name is only a handle for the symbol, NOT a device address: never hand
// `g_scratch` itself to a kernel. The address comes from cudaGetSymbolAddress (scratch_base()).
static __device__ __align__(256) unsigned char g_scratch[CCMPI_NOALLOC_PAD_BYTES + CCMPI_NOALLOC_MAX_BYTES];

namespace {

using namespace gemm;

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_base[kMaxDevices] = {nullptr};

// Device address of g_scratch on the CURRENT device; resolved once per device.
int scratch_base(unsigned char** out) {
    int dev = -1;
    CCMPI_CUDA_CHECK(cudaGetDevice(&dev));
    CCMPI_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_base[dev] == nullptr) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            ccmpi::set_error("cudaGetSymbolAddress(g_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
            return CCMPI_ERR_CUDA;
        }
        g_base[dev] = static_cast<unsigned char*>(ptr);
    }
    *out = g_base[dev];
    return CCMPI_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// bf16 row-major [rows][cols] operand, 128-byte swizzled box; no descriptor cache on this path
int encode(CUtensorMap* out, const void* base, long long rows, long long cols, long long ld, int box_rows, int box_cols) {
    auto fn = ccmpi::drv<EncodeTiledFn>("cuTensorMapEncodeTiled");
    if (!fn) return CCMPI_ERR_CUDA;
    CCMPI_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0,
                  "operand %p (ld %lld) is not 16-byte aligned", base, ld);
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)(ld * 2)};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    const cuuint32_t ones[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, ones,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        ccmpi::set_error("cuTensorMapEncodeTiled failed: CUresult %d", (int)r);
        return CCMPI_ERR_CUDA;
    }
    return CCMPI_OK;
}

template <int BN, bool A_MN, bool B_MN, typename OutT>
int run(const Launch& lp, int grid, cudaStream_t stream) {
    auto kern = gemm_kernel<BN, A_MN, B_MN, OutT, RsNSlotPolicy>;
    CCMPI_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
    kern<<<grid, NUM_THREADS, Cfg<BN>::SMEM_BYTES, stream>>>(lp);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

template <int BN, typename OutT>
int run_layout(const Launch& lp, bool a_km, bool b_kn, int grid, cudaStream_t stream) {
    if (!a_km && !b_kn) return run<BN, false, false, OutT>(lp, grid, stream);
    if (!a_km && b_kn) return run<BN, false, true, OutT>(lp, grid, stream);
    if (a_km && !b_kn) return run<BN, true, false, OutT>(lp, grid, stream);
    return run<BN, true, true, OutT>(lp, grid, stream);
}

}  // namespace

extern "C" size_t ccmpi_gemm_rs_local_noalloc_max_bytes(void) { return CCMPI_NOALLOC_MAX_BYTES; }

extern "C" int ccmpi_gemm_rs_local_noalloc(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                                           int ldc, int a_km, int b_kn, int out_dtype, void* stream) {
    CCMPI_REQUIRE(A && B && C, "null operand");
    CCMPI_REQUIRE(M > 0 && N > 0 && K > 0 && M % 8 == 0 && N % 8 == 0 && K % 8 == 0,
                  "extents must be positive multiples of 8 (M=%d N=%d K=%d)", M, N, K);
    CCMPI_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "leading dimensions must be multiples of 8");
    CCMPI_REQUIRE(out_dtype == CCMPI_BF16 || out_dtype == CCMPI_F32, "output must be bf16 or fp32");
    CCMPI_REQUIRE(N % 128 == 0, "N = %d must be a multiple of 128 (tile width of the scatter epilogue)", N);
    const size_t esz = out_dtype == CCMPI_BF16 ? 2 : 4;
    const size_t need = (size_t)M * (size_t)N * esz;
    if (need > CCMPI_NOALLOC_MAX_BYTES) {
        ccmpi::set_error("problem needs %zu bytes of staging (M=%d x N=%d x %zu B) but the built-in scratch holds %u; "
                         "use ccmpi_fused_gemm_rs with caller-provided memory for larger outputs",
                         need, M, N, esz, (unsigned)CCMPI_NOALLOC_MAX_BYTES);
        return CCMPI_ERR_LIMIT;
    }
    const int BN = (N % 256 == 0) ? 256 : 128;
    const int m_blocks = (M + BM - 1) / BM;
    const int flag_words = m_blocks * (N / BN);
    if (flag_words > CCMPI_PAD_USER_WORDS) {
        ccmpi::set_error("problem needs %d flag words, the built-in scratch holds %d", flag_words, CCMPI_PAD_USER_WORDS);
        return CCMPI_ERR_LIMIT;
    }
    unsigned char* base = nullptr;
    int rc = scratch_base(&base);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // flags of the previous call are still there: they must read as "not arrived" again
    CCMPI_CUDA_CHECK(cudaMemsetAsync(base, 0, CCMPI_NOALLOC_PAD_BYTES, st));

    Launch lp;
    memset(&lp, 0, sizeof(lp));
    lp.nvr = 1;
    RankParams& rp = lp.r[0];
    if (!a_km) rc = encode(&rp.tmap_a, A, M, K, lda, BM, BK);
    else rc = encode(&rp.tmap_a, A, K, M, lda, BK, 64);
    if (rc) return rc;
    rp.tmap_a2 = rp.tmap_a;
    if (!b_kn) rc = encode(&rp.tmap_b, B, N, K, ldb, BN, BK);
    else rc = encode(&rp.tmap_b, B, K, N, ldb, BK, 64);
    if (rc) return rc;
    rp.C = C;
    rp.ldc = ldc;
    rp.M = M; rp.N = N; rp.K = K;
    rp.comm.rank = 0;
    rp.comm.size = 1;
    rp.comm.heap[0] = base + CCMPI_NOALLOC_PAD_BYTES;              // staging starts right after the flags
    rp.comm.pad[0] = reinterpret_cast<uint32_t*>(base);
    rp.comm.mc_heap = nullptr;
    rp.off1 = 0;
    rp.chunk = N;       // one member owns every column
    rp.flag_base = 0;
    rp.epoch = 1;
    rp.aux0 = BN;
    const int sms = ccmpi::device_sm_count();
    const int tiles = m_blocks * (N / BN);
    const int grid = tiles < sms ? tiles : sms;
    if (BN == 256) {
        if (out_dtype == CCMPI_BF16) return run_layout<256, __nv_bfloat16>(lp, a_km != 0, b_kn != 0, grid, st);
        return run_layout<256, float>(lp, a_km != 0, b_kn != 0, grid, st);
    }
    if (out_dtype == CCMPI_BF16) return run_layout<128, __nv_bfloat16>(lp, a_km != 0, b_kn != 0, grid, st);
    return run_layout<128, float>(lp, a_km != 0, b_kn != 0, grid, st);
}
