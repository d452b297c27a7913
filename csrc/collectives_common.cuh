// Shared by the collective kernel files (collectives.cu, collectives_tight.cu): launch parameters, the per-block
// cross-rank barrier, the unrolled grid-stride loop, 16-byte vector reduce helpers and the host-side launch helpers.
// Moved here verbatim from collectives.cu (its SASS is unchanged by the move).
#pragma once
#include "host_common.h"
#include "ptx.cuh"

namespace coll {

constexpr int THREADS = 512;

struct Launch {
    int nvr;
    int pad_[3];
    ccmpi_dev_comm c[CCMPI_MAX_VRANKS];
};

struct Ctx {
    const ccmpi_dev_comm* c;
    int blk;      // block index within the rank
    int nblk;     // blocks per rank
};

__device__ __forceinline__ Ctx make_ctx(const Launch& lp) {
    Ctx x;
    x.c = &lp.c[blockIdx.x % lp.nvr];
    x.blk = blockIdx.x / lp.nvr;
    x.nblk = gridDim.x / lp.nvr;
    return x;
}

// Block-granular barrier across ranks. Block b of rank r publishes epoch e to every member's
// pad word [b][r] and waits until its own pad holds >= e from everyone. Epochs only grow, so
// the flags never need resetting and a kernel can be replayed from a CUDA graph.
__device__ __forceinline__ void block_barrier(const Ctx& x) {
    const ccmpi_dev_comm& c = *x.c;
    __syncthreads();
    uint32_t* self = c.pad[c.rank];
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(self + CCMPI_PAD_EPOCH + x.blk) + 1;
    if (threadIdx.x < c.size) {
        const int peer = threadIdx.x;
        ptx::st_release_sys(c.pad[peer] + CCMPI_PAD_BARRIER + x.blk * CCMPI_MAX_RANKS + c.rank, epoch);
        ptx::wait_flag_ge_sys(self + CCMPI_PAD_BARRIER + x.blk * CCMPI_MAX_RANKS + peer, epoch, peer);
    }
    __syncthreads();
    if (threadIdx.x == 0) *reinterpret_cast<volatile uint32_t*>(self + CCMPI_PAD_EPOCH + x.blk) = epoch;
}

// Grid-stride loop over [lo, hi) that hands the body U indices at a time (i, i+stride, ...), so a
// thread has U independent loads (times the number of peers) in flight: NVLink round trips are
// ~2 us, and bandwidth = bytes in flight / latency.
template <int U, typename Body>
__device__ __forceinline__ void for_range_unrolled(size_t lo, size_t hi, const Ctx& x, Body body) {
    const size_t stride = (size_t)x.nblk * THREADS;
    size_t i = lo + (size_t)x.blk * THREADS + threadIdx.x;
    for (; i + (U - 1) * stride < hi; i += U * stride) body(i, stride, std::integral_constant<int, U>());
    for (; i < hi; i += stride) body(i, stride, std::integral_constant<int, 1>());
}

// ------------------------------------------------------------------------------------------------
// element-wise reduce machinery on 16-byte vectors
// ------------------------------------------------------------------------------------------------
template <int OP, typename A>
__device__ __forceinline__ A apply(A a, A b) {
    if constexpr (OP == CCMPI_SUM) return a + b;
    else if constexpr (OP == CCMPI_MIN) return b < a ? b : a;
    else return b > a ? b : a;
}

// Accumulator type: 16-bit floats accumulate in fp32 and round once.
template <typename T> struct AccOf { using type = T; };
template <> struct AccOf<__nv_bfloat16> { using type = float; };
template <> struct AccOf<__half> { using type = float; };

template <typename T> __device__ __forceinline__ typename AccOf<T>::type to_acc(T v) { return v; }
template <> __device__ __forceinline__ float to_acc<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_acc<__half>(__half v) { return __half2float(v); }

template <typename T, typename A> __device__ __forceinline__ T from_acc(A v) { return static_cast<T>(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16, float>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_acc<__half, float>(float v) { return __float2half_rn(v); }

template <typename T>
struct Vec {
    static constexpr int N = 16 / sizeof(T);
    union {
        uint4 raw;
        T e[N];
    };
    __device__ Vec() {}
};

template <typename T>
struct AccVec {
    using A = typename AccOf<T>::type;
    static constexpr int N = 16 / sizeof(T);
    A e[N];
    __device__ __forceinline__ void init(const uint4& r) {
        Vec<T> v;
        v.raw = r;
#pragma unroll
        for (int i = 0; i < N; ++i) e[i] = to_acc<T>(v.e[i]);
    }
    template <int OP>
    __device__ __forceinline__ void fold(const uint4& r) {
        Vec<T> v;
        v.raw = r;
#pragma unroll
        for (int i = 0; i < N; ++i) e[i] = apply<OP, A>(e[i], to_acc<T>(v.e[i]));
    }
    __device__ __forceinline__ uint4 pack() const {
        Vec<T> v;
#pragma unroll
        for (int i = 0; i < N; ++i) v.e[i] = from_acc<T, A>(e[i]);
        return v.raw;
    }
};

// scale + (optional) cast of a reduced fp32 accumulator vector of N elements, stored to dst
// element offset i0. Only floating outputs take a scale; integer types pass through.
template <typename OutT, typename A, int N>
__device__ __forceinline__ void store_scaled(OutT* dst, size_t i0, const A (&e)[N], float scale) {
    constexpr int OB = N * sizeof(OutT);
    union {
        OutT o[N];
        uint4 q[(OB + 15) / 16];
        uint2 d;
        uint32_t w;
    } u;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        if constexpr (std::is_same<A, float>::value) u.o[i] = from_acc<OutT, float>(e[i] * scale);
        else u.o[i] = static_cast<OutT>(e[i]);
    }
    if constexpr (OB >= 16) {
#pragma unroll
        for (int k = 0; k < OB / 16; ++k) ptx::st_v4(reinterpret_cast<uint8_t*>(dst + i0) + 16 * k, u.q[k]);
    } else if constexpr (OB == 8) {
        *reinterpret_cast<uint2*>(dst + i0) = u.d;
    } else {
        *reinterpret_cast<uint32_t*>(dst + i0) = u.w;
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int fill_launch(Launch& lp, const ccmpi_dev_comm* comms, int nvr) {
    CCMPI_REQUIRE(comms != nullptr && nvr >= 1 && nvr <= CCMPI_MAX_VRANKS, "nvr=%d out of range", nvr);
    memset(&lp, 0, sizeof(lp));
    lp.nvr = nvr;
    for (int i = 0; i < nvr; ++i) {
        lp.c[i] = comms[i];
        CCMPI_REQUIRE(comms[i].size >= 1 && comms[i].size <= CCMPI_MAX_RANKS, "communicator size %d out of range",
                      comms[i].size);
        CCMPI_REQUIRE(comms[i].size == comms[0].size, "emulated ranks disagree on communicator size");
    }
    return CCMPI_OK;
}

// `max_blocks`: 0 = library default; otherwise the cap, optionally OR-ed with CCMPI_BLOCKS_PAIRED:
// the grid is then even and launched as 2-CTA clusters, so the kernel occupies whole TPCs. A
// reduction that overlaps a CTA-pair GEMM whose CTAs wait on peers must not fragment the SM pairs
// that GEMM needs to become fully resident (see parallel/dp.py). A paired count is returned negative.
static int clamp_blocks(size_t work_items, int max_blocks) {
    const bool paired = max_blocks > 0 && (max_blocks & CCMPI_BLOCKS_PAIRED) != 0;
    if (max_blocks > 0) max_blocks &= CCMPI_BLOCKS_PAIRED - 1;
    int cap = max_blocks > 0 ? max_blocks : 128;
    if (cap > CCMPI_MAX_BLOCKS) cap = CCMPI_MAX_BLOCKS;
    size_t want = (work_items + THREADS - 1) / THREADS;
    // a few vectors per thread before adding blocks: small messages are latency-bound and a
    // barrier among few blocks is cheaper; large ones need every SM's worth of loads in flight
    want = (want + 7) / 8;
    if (want < 1) want = 1;
    int blocks = (int)(want < (size_t)cap ? want : (size_t)cap);
    if (!paired) return blocks;
    blocks = (blocks + 1) & ~1;
    if (blocks > cap) blocks = cap & ~1;
    if (blocks < 2) blocks = 2;
    return -blocks;
}

template <typename... Args>
static int launch(void (*kern)(Args...), int nvr, int blocks, cudaStream_t stream, Args... args) {
    void* ptrs[] = {(void*)&args...};
    const bool paired = blocks < 0;
    if (paired) blocks = -blocks;
    if (nvr > 1) {
        // emulated ranks spin on one another: co-residency must be guaranteed, so the grid is
        // clamped to what the device can hold at once (every kernel here is grid-stride)
        int per_sm = 0;
        CCMPI_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)kern, THREADS, 0));
        const int cap = per_sm * ccmpi::device_sm_count() / nvr;
        CCMPI_REQUIRE(cap >= 1, "device cannot hold %d emulated ranks at once", nvr);
        if (blocks > cap) blocks = cap;
        CCMPI_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)kern, dim3(blocks * nvr), dim3(THREADS), ptrs, 0, stream));
    } else if (paired) {
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3(blocks);
        cfg.blockDim = dim3(THREADS);
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        CCMPI_CUDA_CHECK(cudaLaunchKernelExC(&cfg, (const void*)kern, ptrs));
    } else {
        CCMPI_CUDA_CHECK(cudaLaunchKernel((const void*)kern, dim3(blocks), dim3(THREADS), ptrs, 0, stream));
    }
    return CCMPI_OK;
}

static int max_align(size_t v) {
    for (int a = 16; a > 1; a >>= 1)
        if ((v & (size_t)(a - 1)) == 0) return a;
    return 1;
}

}  // namespace coll
