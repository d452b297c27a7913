// Group-size-specialised variants of the two pull-style reductions (opt-in: CCMPI_COLL_TIGHT=1).
//
// The kernels in collectives.cu keep their in-flight vectors in `uint4 r[U][CCMPI_MAX_RANKS]`, sized for 8
// members whatever the communicator size, under the 128-register cap of 512-thread blocks; cuobjdump shows
// part of that array on the stack (312 B for the two-shot all-reduce, 608 B for the reduce-scatter,
// profiles/r1_kernel_resources.txt). Here the array is sized by a template parameter MAXR in {2, 4, 8} and the
// unroll depth is raised so that 16 vectors per thread are in flight at every group size:
//     MAXR = 2: U = 8      MAXR = 4: U = 4      MAXR = 8: U = 2
// Same protocol, same rank-ordered reduction (bit-identical results), same barriers.
//
// Run and measured in round 2 (tests/test_gpu_tight.py; sweeps profiles/r2_sweep_{2,4,8}gpu.jsonl: +3..44 %
// over the generic kernels) and now the default for two-shot all-reduce and pull reduce-scatter; the 2-GPU
// reduce-scatter went from 336 to 602 GB/s at 256 MB (NCCL 390).
#include "collectives_common.cuh"

namespace coll {

template <int MAXR> struct Unroll { static constexpr int value = 16 / MAXR; };

// reduce `U` vectors at element offsets given by `addr(u)` over the first `size` members, rank order
template <typename T, int OP, int MAXR, int U, typename AddrFn, typename StoreFn>
__device__ __forceinline__ void pull_reduce(const ccmpi_dev_comm& c, AddrFn addr, StoreFn store) {
    uint4 r[U][MAXR];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
        for (int p = 0; p < MAXR; ++p)
            if (p < c.size) r[u][p] = ptx::ld_v4(c.heap[p] + addr(u));
#pragma unroll
    for (int u = 0; u < U; ++u) {
        AccVec<T> acc;
        acc.init(r[u][0]);
#pragma unroll
        for (int p = 1; p < MAXR; ++p)
            if (p < c.size) acc.template fold<OP>(r[u][p]);
        store(u, acc);
    }
}

// reduce-scatter (pull), 16-byte vector path of reduce_scatter_kernel<T, OP, true, false>
template <typename T, int OP, int MAXR>
__global__ void __launch_bounds__(THREADS)
reduce_scatter_tight_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t rows, size_t row_elems,
                            size_t src_row_stride, size_t dst_row_stride) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    block_barrier(x);
    uint8_t* dst = c.heap[c.rank] + dst_off;
    const size_t col0 = (size_t)c.rank * row_elems * sizeof(T);
    constexpr int N = 16 / sizeof(T);
    const size_t vpr = row_elems / N;
    const size_t total = rows * vpr;
    for_range_unrolled<Unroll<MAXR>::value>(0, total, x, [&](size_t i0, size_t stride, auto uc) {
        constexpr int U = decltype(uc)::value;
        size_t so[U], dofs[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = i0 + u * stride;
            const size_t row = rows == 1 ? 0 : i / vpr, w = i - row * vpr;
            so[u] = src_off + row * src_row_stride + col0 + w * 16;
            dofs[u] = row * dst_row_stride + w * 16;
        }
        pull_reduce<T, OP, MAXR, U>(c, [&](int u) { return so[u]; },
                                    [&](int u, const AccVec<T>& acc) { ptx::st_v4(dst + dofs[u], acc.pack()); });
    });
    block_barrier(x);
}

// all-reduce, two-shot (allreduce_twoshot_kernel): reduce own slice from all members, park it, gather the slices
template <typename T, int OP, typename OutT, int MAXR>
__global__ void __launch_bounds__(THREADS)
allreduce_twoshot_tight_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t tmp_off, size_t count,
                               float scale) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    using A = typename AccOf<T>::type;
    constexpr int N = 16 / sizeof(T);
    const size_t nvec = count / N;
    const size_t per = (nvec + c.size - 1) / c.size;
    block_barrier(x);
    {
        const size_t lo = per * c.rank, hi = min(nvec, lo + per);
        OutT* tmp = reinterpret_cast<OutT*>(c.heap[c.rank] + tmp_off);
        for_range_unrolled<Unroll<MAXR>::value>(lo, hi, x, [&](size_t i, size_t stride, auto uc) {
            constexpr int U = decltype(uc)::value;
            pull_reduce<T, OP, MAXR, U>(c, [&](int u) { return src_off + (i + u * stride) * 16; },
                                        [&](int u, const AccVec<T>& acc) {
                                            store_scaled<OutT, A, N>(tmp, (i + u * stride) * N, acc.e, scale);
                                        });
        });
    }
    block_barrier(x);
    {
        static_assert(N * sizeof(OutT) == 16, "tight two-shot: output vectors are 16 bytes (no cast)");
        uint8_t* dst = c.heap[c.rank] + dst_off;
        for (int s = 0; s < c.size; ++s) {
            const int p = (c.rank + s) % c.size;
            const size_t lo = per * p, hi = min(nvec, lo + per);
            const uint8_t* from = c.heap[p] + tmp_off;
            for_range_unrolled<8>(lo, hi, x, [&](size_t i, size_t stride, auto uc) {
                constexpr int U = decltype(uc)::value;
                uint4 v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) v[u] = ptx::ld_v4(from + (i + u * stride) * 16);
#pragma unroll
                for (int u = 0; u < U; ++u) ptx::st_v4(dst + (i + u * stride) * 16, v[u]);
            });
        }
    }
    block_barrier(x);
}

template <typename T, int OP>
static int rs_tight(const Launch& lp, int nvr, size_t src_off, size_t dst_off, size_t rows, size_t row_elems, size_t sstride,
                    size_t dstride, int max_blocks, cudaStream_t st) {
    constexpr int N = 16 / sizeof(T);
    const int blocks = clamp_blocks(rows * (row_elems / N), max_blocks);
    const int n = lp.c[0].size;
    if (n <= 2) return launch(reduce_scatter_tight_kernel<T, OP, 2>, nvr, blocks, st, lp, src_off, dst_off, rows, row_elems, sstride, dstride);
    if (n <= 4) return launch(reduce_scatter_tight_kernel<T, OP, 4>, nvr, blocks, st, lp, src_off, dst_off, rows, row_elems, sstride, dstride);
    return launch(reduce_scatter_tight_kernel<T, OP, 8>, nvr, blocks, st, lp, src_off, dst_off, rows, row_elems, sstride, dstride);
}

template <typename T, int OP>
static int ar_tight(const Launch& lp, int nvr, size_t src_off, size_t dst_off, size_t tmp_off, size_t count, float scale,
                    int max_blocks, cudaStream_t st) {
    constexpr int N = 16 / sizeof(T);
    const int n = lp.c[0].size;
    const int blocks = clamp_blocks(count / N / n, max_blocks);
    if (n <= 2) return launch(allreduce_twoshot_tight_kernel<T, OP, T, 2>, nvr, blocks, st, lp, src_off, dst_off, tmp_off, count, scale);
    if (n <= 4) return launch(allreduce_twoshot_tight_kernel<T, OP, T, 4>, nvr, blocks, st, lp, src_off, dst_off, tmp_off, count, scale);
    return launch(allreduce_twoshot_tight_kernel<T, OP, T, 8>, nvr, blocks, st, lp, src_off, dst_off, tmp_off, count, scale);
}

}  // namespace coll

using namespace coll;

// Same contract as ccmpi_reduce_scatter for fp32 / bf16 with 16-byte aligned rows; anything else returns
// CCMPI_ERR_UNSUPPORTED and the caller uses the regular entry point.
extern "C" int ccmpi_reduce_scatter_tight(const ccmpi_dev_comm* comms, int nvr, int dtype, int op, size_t src_off, size_t dst_off,
                                          size_t rows, size_t row_elems, size_t src_row_stride, size_t dst_row_stride,
                                          int max_blocks, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (rows == 0 || row_elems == 0) return CCMPI_OK;
    const size_t es = dtype == CCMPI_F32 ? 4 : 2;
    const bool vec = ((src_off | dst_off | (row_elems * es) | (rows > 1 ? (src_row_stride | dst_row_stride) : 0)) & 15) == 0;
    if ((dtype != CCMPI_F32 && dtype != CCMPI_BF16) || !vec || op < CCMPI_SUM || op > CCMPI_MAX) {
        ccmpi::set_error("tight reduce-scatter: fp32 / bf16 with 16-byte aligned rows only");
        return CCMPI_ERR_UNSUPPORTED;
    }
    cudaStream_t st = (cudaStream_t)stream;
#define CCMPI_RS_TIGHT(T)                                                                                                        \
    switch (op) {                                                                                                                \
        case CCMPI_SUM: return rs_tight<T, CCMPI_SUM>(lp, nvr, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, max_blocks, st); \
        case CCMPI_MIN: return rs_tight<T, CCMPI_MIN>(lp, nvr, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, max_blocks, st); \
        default: return rs_tight<T, CCMPI_MAX>(lp, nvr, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, max_blocks, st);        \
    }
    if (dtype == CCMPI_F32) { CCMPI_RS_TIGHT(float) }
    CCMPI_RS_TIGHT(__nv_bfloat16)
#undef CCMPI_RS_TIGHT
}

// Two-shot all-reduce, same dtype in and out (a scale is applied to float results), fp32 / bf16, SUM / MIN / MAX.
extern "C" int ccmpi_allreduce_twoshot_tight(const ccmpi_dev_comm* comms, int nvr, int dtype, int op, size_t src_off, size_t dst_off,
                                             size_t tmp_off, size_t count, float scale, int max_blocks, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (count == 0) return CCMPI_OK;
    const size_t es = dtype == CCMPI_F32 ? 4 : 2;
    const bool ok = (dtype == CCMPI_F32 || dtype == CCMPI_BF16) && ((src_off | dst_off | tmp_off) & 15) == 0 &&
                    (count * es) % 16 == 0 && op >= CCMPI_SUM && op <= CCMPI_MAX;
    if (!ok) {
        ccmpi::set_error("tight two-shot all-reduce: fp32 / bf16, 16-byte aligned offsets and sizes only");
        return CCMPI_ERR_UNSUPPORTED;
    }
    cudaStream_t st = (cudaStream_t)stream;
#define CCMPI_AR_TIGHT(T)                                                                                        \
    switch (op) {                                                                                                \
        case CCMPI_SUM: return ar_tight<T, CCMPI_SUM>(lp, nvr, src_off, dst_off, tmp_off, count, scale, max_blocks, st); \
        case CCMPI_MIN: return ar_tight<T, CCMPI_MIN>(lp, nvr, src_off, dst_off, tmp_off, count, scale, max_blocks, st); \
        default: return ar_tight<T, CCMPI_MAX>(lp, nvr, src_off, dst_off, tmp_off, count, scale, max_blocks, st);        \
    }
    if (dtype == CCMPI_F32) { CCMPI_AR_TIGHT(float) }
    CCMPI_AR_TIGHT(__nv_bfloat16)
#undef CCMPI_AR_TIGHT
}
