// Stand-alone collective kernels over peer-mapped symmetric memory (NVLink 5 / NVSwitch).
//
// What MPI does for the reference -- MPI_Barrier (comm.py:16), MPI_Allreduce (:22),
// MPI_Allgather (:29), MPI_Reduce_scatter_block (:36), MPI_Alltoall (:61) -- and what the
// reference hand-rolls over Send/Recv (myAllreduce :63-107, myAlltoall :110-159, myAlltoall2
// :162-199) are single kernels here: every rank launches the same grid, block b of every rank
// forms a little group that synchronises through epoch-numbered flags in the ranks' signal
// pads (st.release.sys / ld.acquire.sys), and the payload moves with 16-byte loads/stores on
// peer pointers or with multimem instructions that the NVSwitch executes.
//
// "Virtual ranks": the launch carries up to CCMPI_MAX_VRANKS communicators. In production
// there is one. In single-GPU tests several emulated ranks run inside ONE cooperative launch
// (blockIdx.x % nvr selects the rank), which exercises the same protocol code without ever
// having two kernels wait on each other on one GPU.
#include "collectives_common.cuh"
#include "sr.cuh"

namespace coll {

// ------------------------------------------------------------------------------------------------
// barrier
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS) barrier_kernel(const __grid_constant__ Launch lp) {
    Ctx x = make_ctx(lp);
    block_barrier(x);
}

// ------------------------------------------------------------------------------------------------
// all-reduce, one-shot: every rank pulls all n copies and folds them in rank order 0..n-1,
// the same order as the root loop of myAllreduce (comm.py:85-93), so results are bit-identical
// on every rank and run to run.
// ------------------------------------------------------------------------------------------------
template <typename T, int OP, typename OutT, bool VECTOR>
__global__ void __launch_bounds__(THREADS)
allreduce_oneshot_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t count, float scale) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    using A = typename AccOf<T>::type;
    block_barrier(x);
    OutT* dst = reinterpret_cast<OutT*>(c.heap[c.rank] + dst_off);
    if constexpr (VECTOR) {
        constexpr int N = 16 / sizeof(T);
        const size_t nvec = count / N;
        for (size_t i = (size_t)x.blk * THREADS + threadIdx.x; i < nvec; i += (size_t)x.nblk * THREADS) {
            uint4 r[CCMPI_MAX_RANKS];
#pragma unroll
            for (int p = 0; p < CCMPI_MAX_RANKS; ++p)
                if (p < c.size) r[p] = ptx::ld_v4(c.heap[p] + src_off + i * 16);
            AccVec<T> acc;
            acc.init(r[0]);
#pragma unroll
            for (int p = 1; p < CCMPI_MAX_RANKS; ++p)
                if (p < c.size) acc.template fold<OP>(r[p]);
            store_scaled<OutT, A, N>(dst, i * N, acc.e, scale);
        }
    } else {
        for (size_t i = (size_t)x.blk * THREADS + threadIdx.x; i < count; i += (size_t)x.nblk * THREADS) {
            A acc = to_acc<T>(reinterpret_cast<const T*>(c.heap[0] + src_off)[i]);
            for (int p = 1; p < c.size; ++p)
                acc = apply<OP, A>(acc, to_acc<T>(reinterpret_cast<const T*>(c.heap[p] + src_off)[i]));
            if constexpr (std::is_same<A, float>::value) dst[i] = from_acc<OutT, float>(acc * scale);
            else dst[i] = static_cast<OutT>(acc);
        }
    }
    block_barrier(x);
}

// ------------------------------------------------------------------------------------------------
// all-reduce, two-shot: rank r reduces slice r (pull, rank order), parks it in symmetric
// scratch, then everyone gathers the n reduced slices. 2(n-1)/n of the payload crosses each
// link instead of (n-1).
// ------------------------------------------------------------------------------------------------
template <typename T, int OP, typename OutT>
__global__ void __launch_bounds__(THREADS)
allreduce_twoshot_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t tmp_off,
                         size_t count, float scale) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    using A = typename AccOf<T>::type;
    constexpr int N = 16 / sizeof(T);
    const size_t nvec = count / N;                       // host guarantees divisibility
    const size_t per = (nvec + c.size - 1) / c.size;     // vectors per slice
    block_barrier(x);
    {
        const size_t lo = per * c.rank, hi = min(nvec, lo + per);
        OutT* tmp = reinterpret_cast<OutT*>(c.heap[c.rank] + tmp_off);
        auto reduce_body = [&](size_t i, size_t stride, auto uc) {
            constexpr int U = decltype(uc)::value;
            uint4 r[U][CCMPI_MAX_RANKS];
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int p = 0; p < CCMPI_MAX_RANKS; ++p)
                    if (p < c.size) r[u][p] = ptx::ld_v4(c.heap[p] + src_off + (i + u * stride) * 16);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                AccVec<T> acc;
                acc.init(r[u][0]);
#pragma unroll
                for (int p = 1; p < CCMPI_MAX_RANKS; ++p)
                    if (p < c.size) acc.template fold<OP>(r[u][p]);
                store_scaled<OutT, A, N>(tmp, (i + u * stride) * N, acc.e, scale);
            }
        };
        if (c.size <= 2) for_range_unrolled<4>(lo, hi, x, reduce_body);
        else if (c.size <= 4) for_range_unrolled<2>(lo, hi, x, reduce_body);
        else for_range_unrolled<1>(lo, hi, x, reduce_body);
    }
    block_barrier(x);
    {
        constexpr size_t OV = N * sizeof(OutT);  // bytes of one output vector
        uint8_t* dst = c.heap[c.rank] + dst_off;
        for (int s = 0; s < c.size; ++s) {
            const int p = (c.rank + s) % c.size;  // own slice first, then spread over peers
            const size_t lo = per * p, hi = min(nvec, lo + per);
            const uint8_t* from = c.heap[p] + tmp_off;
            if constexpr (OV == 16) {
                for_range_unrolled<8>(lo, hi, x, [&](size_t i, size_t stride, auto uc) {
                    constexpr int U = decltype(uc)::value;
                    uint4 v[U];
#pragma unroll
                    for (int u = 0; u < U; ++u) v[u] = ptx::ld_v4(from + (i + u * stride) * 16);
#pragma unroll
                    for (int u = 0; u < U; ++u) ptx::st_v4(dst + (i + u * stride) * 16, v[u]);
                });
                continue;
            }
            for (size_t i = lo + (size_t)x.blk * THREADS + threadIdx.x; i < hi; i += (size_t)x.nblk * THREADS) {
                if constexpr (OV == 16) {
                    ptx::st_v4(dst + i * 16, ptx::ld_v4(from + i * 16));
                } else if constexpr (OV == 8) {
                    *reinterpret_cast<uint2*>(dst + i * 8) = ptx::ld_v2(from + i * 8);
                } else {
#pragma unroll
                    for (int k = 0; k < (int)(OV / 16); ++k)
                        ptx::st_v4(dst + i * OV + 16 * k, ptx::ld_v4(from + i * OV + 16 * k));
                }
            }
        }
    }
    block_barrier(x);
}

// ------------------------------------------------------------------------------------------------
// all-reduce, NVLS: the switch reduces (multimem.ld_reduce) and broadcasts (multimem.st).
// Each rank owns slice r; one load + one store per 16 bytes, no per-peer loops at all.
// ------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ uint4 mc_ld_reduce(const void* p);
template <> __device__ __forceinline__ uint4 mc_ld_reduce<float>(const void* p) { return ptx::multimem_ld_reduce_add_f32x4(p); }
template <> __device__ __forceinline__ uint4 mc_ld_reduce<__nv_bfloat16>(const void* p) { return ptx::multimem_ld_reduce_add_bf16x8(p); }
template <> __device__ __forceinline__ uint4 mc_ld_reduce<__half>(const void* p) { return ptx::multimem_ld_reduce_add_f16x8(p); }

template <typename T, typename OutT>
__global__ void __launch_bounds__(THREADS)
allreduce_nvls_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t count, float scale) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    constexpr int N = 16 / sizeof(T);
    constexpr int G = (sizeof(OutT) < sizeof(T)) ? (int)(sizeof(T) / sizeof(OutT)) : 1;  // input vectors per 16-byte output
    const size_t ngrp = count / (N * G);
    const size_t per = (ngrp + c.size - 1) / c.size;
    const size_t lo = per * c.rank, hi = min(ngrp, lo + per);
    const uint8_t* msrc = c.mc_heap + src_off;
    uint8_t* mdst = c.mc_heap + dst_off;
    const bool identity = (scale == 1.0f) && std::is_same<T, OutT>::value;
    block_barrier(x);
    if constexpr (std::is_same<T, OutT>::value) {
        if (identity) {
            for_range_unrolled<8>(lo, hi, x, [&](size_t g, size_t stride, auto uc) {
                constexpr int U = decltype(uc)::value;
                uint4 v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) v[u] = mc_ld_reduce<T>(msrc + (g + u * stride) * 16);
#pragma unroll
                for (int u = 0; u < U; ++u) ptx::multimem_st_v4(mdst + (g + u * stride) * 16, v[u]);
            });
            block_barrier(x);
            return;
        }
    }
    if constexpr (std::is_same<T, OutT>::value) {
        // scaled, same type (the data-parallel gradient path: scale = 1/dp): eight reductions in flight per
        // thread, like the unscaled path -- one at a time leaves the switch round trip exposed
        for_range_unrolled<8>(lo, hi, x, [&](size_t g, size_t stride, auto uc) {
            constexpr int U = decltype(uc)::value;
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = mc_ld_reduce<T>(msrc + (g + u * stride) * 16);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                AccVec<T> a;
                a.init(v[u]);
#pragma unroll
                for (int i = 0; i < N; ++i) a.e[i] = a.e[i] * scale;
                ptx::multimem_st_v4(mdst + (g + u * stride) * 16, a.pack());
            }
        });
        block_barrier(x);
        return;
    }
    for (size_t g = lo + (size_t)x.blk * THREADS + threadIdx.x; g < hi; g += (size_t)x.nblk * THREADS) {
        if constexpr (std::is_same<T, OutT>::value) {
            // (handled above)
        } else if constexpr (sizeof(OutT) < sizeof(T)) {
            // fp32 in -> 16-bit out: two input vectors make one output vector
            Vec<OutT> o;
#pragma unroll
            for (int k = 0; k < G; ++k) {
                Vec<T> v;
                v.raw = mc_ld_reduce<T>(msrc + (g * G + k) * 16);
#pragma unroll
                for (int i = 0; i < N; ++i) o.e[k * N + i] = from_acc<OutT, float>(to_acc<T>(v.e[i]) * scale);
            }
            ptx::multimem_st_v4(mdst + g * 16, o.raw);
        } else {
            // 16-bit in -> fp32 out: one input vector makes two output vectors
            Vec<T> v;
            v.raw = mc_ld_reduce<T>(msrc + g * 16);
            constexpr int R = sizeof(OutT) / sizeof(T);
#pragma unroll
            for (int k = 0; k < R; ++k) {
                Vec<OutT> o;
#pragma unroll
                for (int i = 0; i < N / R; ++i) o.e[i] = static_cast<OutT>(to_acc<T>(v.e[k * (N / R) + i]) * scale);
                ptx::multimem_st_v4(mdst + (g * R + k) * 16, o.raw);
            }
        }
    }
    block_barrier(x);
}

// ------------------------------------------------------------------------------------------------
// Data-parallel gradient all-reduce FUSED WITH THE OPTIMIZER (bf16, NVLS). Rank r owns slice r of the
// bucket: it lets the switch sum that slice of every replica's gradient (multimem.ld_reduce), applies
//     w = SR(w - lr_eff * sum)            (fp32 math, stochastic rounding, sr.cuh; lr_eff = lr / dp)
// to ITS copy of the weights and multicasts the new weights into every replica (multimem.st). The
// reduced gradient is never written anywhere, there is no optimizer pass, and since each weight is
// computed by exactly one rank the replicas stay bit-identical. Weights and gradients live at the same
// symmetric offsets on all members (parallel/dp.py lays both out as one flat buffer each).
// Per parameter and rank: 2 B read over NVLink (its slice only) + 2 B local read + 2 B multicast write.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS)
allreduce_sgd_nvls_kernel(const __grid_constant__ Launch lp, size_t grad_off, size_t w_off, size_t count, float lr_eff,
                          unsigned long long base_index, uint32_t step, uint32_t seed) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    const size_t nvec = count / 8;
    const size_t per = (nvec + c.size - 1) / c.size;
    const size_t lo = per * c.rank, hi = min(nvec, lo + per);
    const uint8_t* mgrad = c.mc_heap + grad_off;
    uint8_t* mw = c.mc_heap + w_off;
    const uint8_t* w_loc = c.heap[c.rank] + w_off;
    block_barrier(x);       // every replica's gradients are final, nobody still reads the old weights
    for_range_unrolled<8>(lo, hi, x, [&](size_t g, size_t stride, auto uc) {
        constexpr int U = decltype(uc)::value;
        uint4 gs[U], ws[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            gs[u] = ptx::multimem_ld_reduce_add_bf16x8(mgrad + (g + u * stride) * 16);
            ws[u] = ptx::ld_v4(w_loc + (g + u * stride) * 16);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&gs[u]);
            const __nv_bfloat162* wh = reinterpret_cast<const __nv_bfloat162*>(&ws[u]);
            uint4 out;
            uint32_t* o = reinterpret_cast<uint32_t*>(&out);
            const unsigned long long pair0 = (base_index >> 1) + (unsigned long long)(g + u * stride) * 4;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 gf = __bfloat1622float2(gh[e]);
                const float2 wf = __bfloat1622float2(wh[e]);
                o[e] = sr::pack2(sr::axpy(wf.x, lr_eff, gf.x), sr::axpy(wf.y, lr_eff, gf.y), sr::bits(pair0 + e, step, seed));
            }
            ptx::multimem_st_v4(mw + (g + u * stride) * 16, out);
        }
    });
    block_barrier(x);       // all slices of the new weights have landed everywhere
}

// ------------------------------------------------------------------------------------------------
// all-gather (push): row i of this rank goes to every member's dst at
//     dst + i*dst_row_stride + rank*row_bytes.
// rows = 1 is MPI_Allgather; rows > 1 is the last-axis gather of func_impl.py:89-90 without the
// np.concatenate copy. W = access width in bytes (16 when everything is 16-byte aligned).
// ------------------------------------------------------------------------------------------------
template <int W> struct Word;
template <> struct Word<16> { using type = uint4; };
template <> struct Word<8> { using type = uint2; };
template <> struct Word<4> { using type = uint32_t; };
template <> struct Word<2> { using type = uint16_t; };
template <> struct Word<1> { using type = uint8_t; };

template <int W, bool MC>
__global__ void __launch_bounds__(THREADS)
allgather_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t rows, size_t row_bytes,
                 size_t src_row_stride, size_t dst_row_stride) {
    using WT = typename Word<W>::type;
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    const size_t wpr = row_bytes / W;  // words per row
    const size_t total = rows * wpr;
    block_barrier(x);  // every member's dst is free to be overwritten
    const uint8_t* src = c.heap[c.rank] + src_off;
    const size_t col0 = (size_t)c.rank * row_bytes;
    for_range_unrolled<4>(0, total, x, [&](size_t i0, size_t stride, auto uc) {
        constexpr int U = decltype(uc)::value;
        WT v[U];
        size_t d[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t i = i0 + u * stride;
            const size_t row = rows == 1 ? 0 : i / wpr, w = i - row * wpr;       // flat case: no 64-bit divide
            v[u] = *reinterpret_cast<const WT*>(src + row * src_row_stride + w * W);
            d[u] = dst_off + row * dst_row_stride + col0 + w * W;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if constexpr (MC && W == 16) {
                ptx::multimem_st_v4(c.mc_heap + d[u], v[u]);
            } else {
#pragma unroll
                for (int s = 0; s < CCMPI_MAX_RANKS; ++s) {
                    if (s < c.size) {
                        const int p = (c.rank + s) % c.size;
                        *reinterpret_cast<WT*>(c.heap[p] + d[u]) = v[u];
                    }
                }
            }
        }
    });
    block_barrier(x);  // all pushes have landed everywhere
}

// ------------------------------------------------------------------------------------------------
// reduce-scatter (pull): rank r reduces column block r of every row over all members.
// rows = 1 is MPI_Reduce_scatter_block; rows > 1 is func_impl.py:182-186 (split on the last
// axis, all-to-all, sum) in one pass and in the same rank order as np.sum(axis=0).
// ------------------------------------------------------------------------------------------------
template <typename T, int OP, bool VECTOR, bool MC>
__global__ void __launch_bounds__(THREADS)
reduce_scatter_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t rows, size_t row_elems,
                      size_t src_row_stride, size_t dst_row_stride) {
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    using A = typename AccOf<T>::type;
    block_barrier(x);
    uint8_t* dst = c.heap[c.rank] + dst_off;
    const size_t col0 = (size_t)c.rank * row_elems * sizeof(T);
    if constexpr (VECTOR) {
        constexpr int N = 16 / sizeof(T);
        const size_t vpr = row_elems / N;
        const size_t total = rows * vpr;
        auto body = [&](size_t i0, size_t stride, auto uc) {
            constexpr int U = decltype(uc)::value;
            size_t so[U], dofs[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                const size_t row = rows == 1 ? 0 : i / vpr, w = i - row * vpr;   // flat case: no 64-bit divide
                so[u] = src_off + row * src_row_stride + col0 + w * 16;
                dofs[u] = row * dst_row_stride + w * 16;
            }
            if constexpr (MC) {
                uint4 out[U];
#pragma unroll
                for (int u = 0; u < U; ++u) out[u] = mc_ld_reduce<T>(c.mc_heap + so[u]);
#pragma unroll
                for (int u = 0; u < U; ++u) ptx::st_v4(dst + dofs[u], out[u]);
            } else {
                uint4 r[U][CCMPI_MAX_RANKS];
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int p = 0; p < CCMPI_MAX_RANKS; ++p)
                        if (p < c.size) r[u][p] = ptx::ld_v4(c.heap[p] + so[u]);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    AccVec<T> acc;
                    acc.init(r[u][0]);
#pragma unroll
                    for (int p = 1; p < CCMPI_MAX_RANKS; ++p)
                        if (p < c.size) acc.template fold<OP>(r[u][p]);
                    ptx::st_v4(dst + dofs[u], acc.pack());
                }
            }
        };
        if (MC) for_range_unrolled<8>(0, total, x, body);
        else if (c.size <= 2) for_range_unrolled<4>(0, total, x, body);   // 8 would spill: r[U][8] registers
        else if (c.size <= 4) for_range_unrolled<2>(0, total, x, body);
        else for_range_unrolled<1>(0, total, x, body);
    } else {
        const size_t total = rows * row_elems;
        for (size_t i = (size_t)x.blk * THREADS + threadIdx.x; i < total; i += (size_t)x.nblk * THREADS) {
            const size_t row = i / row_elems, w = i - row * row_elems;
            const size_t s = src_off + row * src_row_stride + col0 + w * sizeof(T);
            A acc = to_acc<T>(*reinterpret_cast<const T*>(c.heap[0] + s));
            for (int p = 1; p < c.size; ++p) acc = apply<OP, A>(acc, to_acc<T>(*reinterpret_cast<const T*>(c.heap[p] + s)));
            *reinterpret_cast<T*>(dst + row * dst_row_stride + w * sizeof(T)) = from_acc<T, A>(acc);
        }
    }
    block_barrier(x);  // nobody may overwrite src before every member has read it
}

// ------------------------------------------------------------------------------------------------
// all-to-all (push): segment j of src lands in member j's dst at segment index `rank`. All
// n-1 peer transfers are in flight from one kernel (what Isend/Irecv + Waitall buy myAlltoall,
// comm.py:136-150) and land directly in dst (no receive buffers, comm.py:152-155). In
// `staggered` mode a block walks peers in (rank+i)%n order like the Sendrecv loop of
// myAlltoall2; otherwise work is interleaved over all peers at vector granularity.
// ------------------------------------------------------------------------------------------------
template <int W>
__global__ void __launch_bounds__(THREADS)
alltoall_kernel(const __grid_constant__ Launch lp, size_t src_off, size_t dst_off, size_t seg_bytes, int staggered) {
    using WT = typename Word<W>::type;
    Ctx x = make_ctx(lp);
    const ccmpi_dev_comm& c = *x.c;
    const size_t wps = seg_bytes / W;
    block_barrier(x);
    const uint8_t* src = c.heap[c.rank] + src_off;
    if (staggered) {
        for (int s = 0; s < c.size; ++s) {
            const int p = (c.rank + s) % c.size;
            const WT* from = reinterpret_cast<const WT*>(src + (size_t)p * seg_bytes);
            WT* to = reinterpret_cast<WT*>(c.heap[p] + dst_off + (size_t)c.rank * seg_bytes);
            for (size_t i = (size_t)x.blk * THREADS + threadIdx.x; i < wps; i += (size_t)x.nblk * THREADS) to[i] = from[i];
        }
    } else {
        const size_t total = wps * c.size;
        // peer-major within a warp-sized run keeps both the local loads and the remote stores
        // coalesced; consecutive runs rotate over the peers so every link is busy at once
        const size_t runs_per_seg = (wps + 31) / 32;
        for_range_unrolled<4>(0, runs_per_seg * 32 * c.size, x, [&](size_t i0, size_t stride, auto uc) {
            constexpr int U = decltype(uc)::value;
            WT v[U];
            size_t w[U];
            int p[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                const size_t run = i / 32;
                p[u] = (int)((run + c.rank) % c.size);
                w[u] = (run / c.size) * 32 + (i & 31);
                if (w[u] < wps) v[u] = reinterpret_cast<const WT*>(src + (size_t)p[u] * seg_bytes)[w[u]];
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (w[u] < wps) reinterpret_cast<WT*>(c.heap[p[u]] + dst_off + (size_t)c.rank * seg_bytes)[w[u]] = v[u];
        });
    }
    block_barrier(x);
}

// ------------------------------------------------------------------------------------------------
// point-to-point: put-with-signal into the peer's segment, and the matching wait. Backs the
// Send/Recv/Isend/Irecv/Sendrecv surface of the mpi4py front end (comm.py:86,99,104-105,140,147,
// 192). Flags are counters in the user area of the signal pad: the sender adds 1, the receiver
// waits for its private expected count (kept in the word after the flag).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS)
put_signal_kernel(const __grid_constant__ Launch lp, int peer, size_t src_off, size_t dst_off, size_t bytes, int flag) {
    const ccmpi_dev_comm& c = lp.c[0];
    const uint8_t* src = c.heap[c.rank] + src_off;
    uint8_t* dst = c.heap[peer] + dst_off;
    if (((src_off | dst_off | bytes) & 15) == 0) {
        for (size_t i = threadIdx.x; i < bytes / 16; i += THREADS) ptx::st_v4(dst + i * 16, ptx::ld_v4(src + i * 16));
    } else {
        for (size_t i = threadIdx.x; i < bytes; i += THREADS) dst[i] = src[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) ptx::red_release_sys_add(c.pad[peer] + CCMPI_PAD_USER + flag, 1u);
}

__global__ void wait_signal_kernel(const __grid_constant__ Launch lp, int flag) {
    const ccmpi_dev_comm& c = lp.c[0];
    uint32_t* f = c.pad[c.rank] + CCMPI_PAD_USER + flag;
    const uint32_t want = *reinterpret_cast<volatile uint32_t*>(f + 1) + 1;
    ptx::wait_flag_ge_sys(f, want, flag);
    *reinterpret_cast<volatile uint32_t*>(f + 1) = want;
}

}  // namespace coll

using namespace coll;

extern "C" int ccmpi_barrier(const ccmpi_dev_comm* comms, int nvr, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    return launch(barrier_kernel, nvr, 1, (cudaStream_t)stream, lp);
}

namespace coll {

template <typename T, int OP>
static int allreduce_typed(const Launch& lp, int nvr, int algo, size_t src_off, size_t dst_off, size_t tmp_off,
                           size_t count, float scale, int out_dtype, int dtype, int max_blocks, cudaStream_t st) {
    constexpr int N = 16 / sizeof(T);
    const bool aligned = ((src_off | dst_off | tmp_off) & 15) == 0;
    const int n = lp.c[0].size;
    constexpr bool is_float = std::is_same<T, float>::value || std::is_same<T, __nv_bfloat16>::value || std::is_same<T, __half>::value;
    const bool cast = out_dtype != dtype;
    CCMPI_REQUIRE(!cast || is_float, "all-reduce output cast is only defined for fp32/bf16/fp16");
    if (algo == CCMPI_AR_NVLS) {
        if constexpr (is_float && OP == CCMPI_SUM) {
            CCMPI_REQUIRE(lp.c[0].mc_heap != nullptr, "NVLS all-reduce needs a multicast mapping");
            CCMPI_REQUIRE(aligned && count % (N * 2) == 0, "NVLS all-reduce needs 16-byte aligned offsets and count %% %d == 0", N * 2);
            const int blocks = clamp_blocks(count / N / n, max_blocks);
            if (!cast) return launch(allreduce_nvls_kernel<T, T>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
            if constexpr (std::is_same<T, float>::value) {
                if (out_dtype == CCMPI_BF16) return launch(allreduce_nvls_kernel<float, __nv_bfloat16>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
            }
            if constexpr (std::is_same<T, __nv_bfloat16>::value) {
                if (out_dtype == CCMPI_F32) return launch(allreduce_nvls_kernel<__nv_bfloat16, float>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
            }
            ccmpi::set_error("NVLS all-reduce: unsupported cast %d -> %d", dtype, out_dtype);
            return CCMPI_ERR_UNSUPPORTED;
        } else {
            ccmpi::set_error("NVLS all-reduce supports SUM over fp32/bf16/fp16 only");
            return CCMPI_ERR_UNSUPPORTED;
        }
    }
    if (algo == CCMPI_AR_TWOSHOT) {
        CCMPI_REQUIRE(aligned && count % N == 0, "two-shot all-reduce needs 16-byte aligned offsets and whole vectors");
        const int blocks = clamp_blocks(count / N / n, max_blocks);
        if (!cast) return launch(allreduce_twoshot_kernel<T, OP, T>, nvr, blocks, st, lp, src_off, dst_off, tmp_off, count, scale);
        if constexpr (std::is_same<T, float>::value) {
            if (out_dtype == CCMPI_BF16) return launch(allreduce_twoshot_kernel<float, OP, __nv_bfloat16>, nvr, blocks, st, lp, src_off, dst_off, tmp_off, count, scale);
        }
        if constexpr (std::is_same<T, __nv_bfloat16>::value) {
            if (out_dtype == CCMPI_F32) return launch(allreduce_twoshot_kernel<__nv_bfloat16, OP, float>, nvr, blocks, st, lp, src_off, dst_off, tmp_off, count, scale);
        }
        ccmpi::set_error("two-shot all-reduce: unsupported cast %d -> %d", dtype, out_dtype);
        return CCMPI_ERR_UNSUPPORTED;
    }
    // one-shot
    const bool vec = aligned && count % N == 0;
    const int blocks = clamp_blocks(vec ? count / N : count, max_blocks);
    if (!cast) {
        if (vec) return launch(allreduce_oneshot_kernel<T, OP, T, true>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
        return launch(allreduce_oneshot_kernel<T, OP, T, false>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
    }
    if constexpr (std::is_same<T, float>::value) {
        if (out_dtype == CCMPI_BF16) {
            if (vec) return launch(allreduce_oneshot_kernel<float, OP, __nv_bfloat16, true>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
            return launch(allreduce_oneshot_kernel<float, OP, __nv_bfloat16, false>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
        }
    }
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        if (out_dtype == CCMPI_F32) {
            if (vec) return launch(allreduce_oneshot_kernel<__nv_bfloat16, OP, float, true>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
            return launch(allreduce_oneshot_kernel<__nv_bfloat16, OP, float, false>, nvr, blocks, st, lp, src_off, dst_off, count, scale);
        }
    }
    ccmpi::set_error("one-shot all-reduce: unsupported cast %d -> %d", dtype, out_dtype);
    return CCMPI_ERR_UNSUPPORTED;
}

template <typename T>
static int allreduce_op(const Launch& lp, int nvr, int algo, int op, size_t src_off, size_t dst_off, size_t tmp_off,
                        size_t count, float scale, int out_dtype, int dtype, int max_blocks, cudaStream_t st) {
    switch (op) {
        case CCMPI_SUM: return allreduce_typed<T, CCMPI_SUM>(lp, nvr, algo, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_MIN: return allreduce_typed<T, CCMPI_MIN>(lp, nvr, algo, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_MAX: return allreduce_typed<T, CCMPI_MAX>(lp, nvr, algo, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
    }
    // mirrors the NotImplementedError of comm.py:94-95, but raised on every rank before any
    // communication so nobody is left waiting
    ccmpi::set_error("reduce op %d is not supported (SUM, MIN, MAX only)", op);
    return CCMPI_ERR_UNSUPPORTED;
}

}  // namespace coll

extern "C" int ccmpi_allreduce(const ccmpi_dev_comm* comms, int nvr, int algo, int dtype, int op, size_t src_off,
                               size_t dst_off, size_t tmp_off, size_t count, float scale, int out_dtype, int max_blocks,
                               void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (count == 0) return CCMPI_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (out_dtype < 0) out_dtype = dtype;
    CCMPI_REQUIRE(dtype >= 0 && dtype <= CCMPI_I64, "all-reduce: unsupported dtype %d", dtype);
    if (algo == CCMPI_AR_AUTO) {
        static const size_t esz[] = {4, 2, 2, 8, 4, 8, 1};
        const size_t bytes = count * esz[dtype];
        const bool nvls_ok = lp.c[0].mc_heap && op == CCMPI_SUM && (dtype == CCMPI_F32 || dtype == CCMPI_BF16 || dtype == CCMPI_F16) &&
                             ((src_off | dst_off) & 15) == 0 && (bytes % 32) == 0;
        const bool vec_ok = ((src_off | dst_off | tmp_off) & 15) == 0 && bytes % 16 == 0;
        // measured (profiles/r2_sweep_2gpu.jsonl, r1_sweep_8gpu.jsonl): one-shot up to 256 KB; between two
        // GPUs NVLS wins up to ~8 MB and the P2P two-shot above; larger groups use NVLS throughout
        if (bytes <= 256 * 1024 || !vec_ok) algo = CCMPI_AR_ONESHOT;
        else if (nvls_ok && (lp.c[0].size > 2 || bytes <= (8u << 20))) algo = CCMPI_AR_NVLS;
        else algo = CCMPI_AR_TWOSHOT;
    }
    if (algo == CCMPI_AR_TWOSHOT && out_dtype == dtype) {
        // group-size-specialised kernel (collectives_tight.cu) whenever it applies
        rc = ccmpi_allreduce_twoshot_tight(comms, nvr, dtype, op, src_off, dst_off, tmp_off, count, scale, max_blocks, stream);
        if (rc != CCMPI_ERR_UNSUPPORTED) return rc;
    }
    if (algo == CCMPI_AR_TWOSHOT_GENERIC) algo = CCMPI_AR_TWOSHOT;
    switch (dtype) {
        case CCMPI_F32: return allreduce_op<float>(lp, nvr, algo, op, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_BF16: return allreduce_op<__nv_bfloat16>(lp, nvr, algo, op, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_F16: return allreduce_op<__half>(lp, nvr, algo, op, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_F64: return allreduce_op<double>(lp, nvr, algo, op, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_I32: return allreduce_op<int32_t>(lp, nvr, algo, op, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
        case CCMPI_I64: return allreduce_op<long long>(lp, nvr, algo, op, src_off, dst_off, tmp_off, count, scale, out_dtype, dtype, max_blocks, st);
    }
    ccmpi::set_error("all-reduce: unsupported dtype %d", dtype);
    return CCMPI_ERR_UNSUPPORTED;
}

// Fused data-parallel step for one bucket: weights[w_off ...] = SR(weights - lr_eff * sum over replicas of
// grads[grad_off ...]), bf16, `count` elements (multiple of 16), both offsets symmetric and 16-byte aligned.
// Needs a multicast mapping (NVLS). lr_eff already contains the 1/dp of the average.
extern "C" int ccmpi_allreduce_sgd_nvls(const ccmpi_dev_comm* comms, int nvr, size_t grad_off, size_t w_off, size_t count,
                                        float lr_eff, unsigned long long base_index, unsigned step, unsigned seed,
                                        int max_blocks, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (count == 0) return CCMPI_OK;
    CCMPI_REQUIRE(lp.c[0].mc_heap != nullptr, "the fused all-reduce + SGD kernel needs a multicast mapping (NVLS)");
    CCMPI_REQUIRE(((grad_off | w_off) & 15) == 0 && count % 16 == 0 && base_index % 2 == 0,
                  "all-reduce + SGD: 16-byte aligned offsets, count %% 16 == 0 (count = %zu), even base index", count);
    const int blocks = clamp_blocks(count / 8 / lp.c[0].size, max_blocks);
    return launch(allreduce_sgd_nvls_kernel, nvr, blocks, (cudaStream_t)stream, lp, grad_off, w_off, count, lr_eff, base_index,
                  (uint32_t)step, (uint32_t)seed);
}

extern "C" int ccmpi_allgather(const ccmpi_dev_comm* comms, int nvr, size_t src_off, size_t dst_off, size_t rows,
                               size_t row_bytes, size_t src_row_stride, size_t dst_row_stride, int use_mc,
                               int max_blocks, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (rows == 0 || row_bytes == 0) return CCMPI_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int W = max_align(src_off | dst_off | row_bytes | (rows > 1 ? (src_row_stride | dst_row_stride) : 0));
    const int blocks = clamp_blocks(rows * (row_bytes / W), max_blocks);
    // multimem.st (one copy leaves the GPU, the switch replicates it) wins up to a few MB; above that the
    // per-peer stores are faster (profiles/r2_sweep_4gpu.jsonl: 569 vs 466 GB/s at 64 MB)
    const bool mc = use_mc && lp.c[0].mc_heap != nullptr && W == 16 && (use_mc == 2 || rows * row_bytes * lp.c[0].size < (8u << 20));
    switch (W) {
        case 16:
            if (mc) return launch(allgather_kernel<16, true>, nvr, blocks, st, lp, src_off, dst_off, rows, row_bytes, src_row_stride, dst_row_stride);
            return launch(allgather_kernel<16, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_bytes, src_row_stride, dst_row_stride);
        case 8: return launch(allgather_kernel<8, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_bytes, src_row_stride, dst_row_stride);
        case 4: return launch(allgather_kernel<4, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_bytes, src_row_stride, dst_row_stride);
        case 2: return launch(allgather_kernel<2, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_bytes, src_row_stride, dst_row_stride);
        default: return launch(allgather_kernel<1, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_bytes, src_row_stride, dst_row_stride);
    }
}

namespace coll {
template <typename T, int OP>
static int rs_typed(const Launch& lp, int nvr, size_t src_off, size_t dst_off, size_t rows, size_t row_elems,
                    size_t sstride, size_t dstride, int use_mc, int max_blocks, cudaStream_t st) {
    constexpr int N = 16 / sizeof(T);
    const size_t row_bytes = row_elems * sizeof(T);
    const bool vec = ((src_off | dst_off | row_bytes | (rows > 1 ? (sstride | dstride) : 0)) & 15) == 0;
    const int blocks = clamp_blocks(vec ? rows * (row_elems / N) : rows * row_elems, max_blocks);
    constexpr bool mc_type = (std::is_same<T, float>::value || std::is_same<T, __nv_bfloat16>::value || std::is_same<T, __half>::value) && OP == CCMPI_SUM;
    if constexpr (mc_type) {
        if (vec && use_mc && lp.c[0].mc_heap)
            return launch(reduce_scatter_kernel<T, OP, true, true>, nvr, blocks, st, lp, src_off, dst_off, rows, row_elems, sstride, dstride);
    }
    if (vec) return launch(reduce_scatter_kernel<T, OP, true, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_elems, sstride, dstride);
    return launch(reduce_scatter_kernel<T, OP, false, false>, nvr, blocks, st, lp, src_off, dst_off, rows, row_elems, sstride, dstride);
}
template <typename T>
static int rs_op(const Launch& lp, int nvr, int op, size_t src_off, size_t dst_off, size_t rows, size_t row_elems,
                 size_t sstride, size_t dstride, int use_mc, int max_blocks, cudaStream_t st) {
    switch (op) {
        case CCMPI_SUM: return rs_typed<T, CCMPI_SUM>(lp, nvr, src_off, dst_off, rows, row_elems, sstride, dstride, use_mc, max_blocks, st);
        case CCMPI_MIN: return rs_typed<T, CCMPI_MIN>(lp, nvr, src_off, dst_off, rows, row_elems, sstride, dstride, use_mc, max_blocks, st);
        case CCMPI_MAX: return rs_typed<T, CCMPI_MAX>(lp, nvr, src_off, dst_off, rows, row_elems, sstride, dstride, use_mc, max_blocks, st);
    }
    ccmpi::set_error("reduce op %d is not supported (SUM, MIN, MAX only)", op);
    return CCMPI_ERR_UNSUPPORTED;
}
}  // namespace coll

extern "C" int ccmpi_reduce_scatter(const ccmpi_dev_comm* comms, int nvr, int dtype, int op, size_t src_off,
                                    size_t dst_off, size_t rows, size_t row_elems, size_t src_row_stride,
                                    size_t dst_row_stride, int use_mc, int max_blocks, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (rows == 0 || row_elems == 0) return CCMPI_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // use_mc: 0 = peer loads, 1 = the library's choice (NVLS for small messages when the group has a multicast mapping),
    // 3 = NVLS at any size, 2 = peer loads with the kernel
    // sized for any group (benchmarks). 0 -- and 1 without NVLS or between two GPUs, where peer loads measured
    // faster -- take the group-size-specialised kernel when dtype / alignment allow (same bits, 1.2-1.7x).
    // NVLS (multimem.ld_reduce) wins while the message is latency-bound; from ~8 MB up the peer-load kernel is
    // faster at every group size measured (profiles/r2_sweep_{2,4}gpu.jsonl: 537 vs 455 GB/s at 64 MB on 4 GPUs)
    const size_t total_bytes = rows * row_elems * (dtype == CCMPI_F64 || dtype == CCMPI_I64 ? 8 : dtype == CCMPI_F32 || dtype == CCMPI_I32 ? 4 : 2) *
                               (size_t)lp.c[0].size;
    if (use_mc == 0 || (use_mc == 1 && (lp.c[0].mc_heap == nullptr || lp.c[0].size <= 2 || total_bytes >= (8u << 20)))) {
        rc = ccmpi_reduce_scatter_tight(comms, nvr, dtype, op, src_off, dst_off, rows, row_elems, src_row_stride,
                                        dst_row_stride, max_blocks, stream);
        if (rc != CCMPI_ERR_UNSUPPORTED) return rc;
        use_mc = 0;
    }
    if (use_mc == 2) use_mc = 0;
    if (use_mc == 3) use_mc = 1;      // NVLS at any size (benchmarks)
    switch (dtype) {
        case CCMPI_F32: return rs_op<float>(lp, nvr, op, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, use_mc, max_blocks, st);
        case CCMPI_BF16: return rs_op<__nv_bfloat16>(lp, nvr, op, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, use_mc, max_blocks, st);
        case CCMPI_F16: return rs_op<__half>(lp, nvr, op, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, use_mc, max_blocks, st);
        case CCMPI_F64: return rs_op<double>(lp, nvr, op, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, use_mc, max_blocks, st);
        case CCMPI_I32: return rs_op<int32_t>(lp, nvr, op, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, use_mc, max_blocks, st);
        case CCMPI_I64: return rs_op<long long>(lp, nvr, op, src_off, dst_off, rows, row_elems, src_row_stride, dst_row_stride, use_mc, max_blocks, st);
    }
    ccmpi::set_error("reduce-scatter: unsupported dtype %d", dtype);
    return CCMPI_ERR_UNSUPPORTED;
}

extern "C" int ccmpi_alltoall(const ccmpi_dev_comm* comms, int nvr, size_t src_off, size_t dst_off, size_t seg_bytes,
                              int staggered, int max_blocks, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, nvr);
    if (rc) return rc;
    if (seg_bytes == 0) return CCMPI_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int W = max_align(src_off | dst_off | seg_bytes);
    const int blocks = clamp_blocks((seg_bytes / W) * lp.c[0].size, max_blocks);
    switch (W) {
        case 16: return launch(alltoall_kernel<16>, nvr, blocks, st, lp, src_off, dst_off, seg_bytes, staggered);
        case 8: return launch(alltoall_kernel<8>, nvr, blocks, st, lp, src_off, dst_off, seg_bytes, staggered);
        case 4: return launch(alltoall_kernel<4>, nvr, blocks, st, lp, src_off, dst_off, seg_bytes, staggered);
        case 2: return launch(alltoall_kernel<2>, nvr, blocks, st, lp, src_off, dst_off, seg_bytes, staggered);
        default: return launch(alltoall_kernel<1>, nvr, blocks, st, lp, src_off, dst_off, seg_bytes, staggered);
    }
}

extern "C" int ccmpi_put_signal(const ccmpi_dev_comm* comms, int nvr, int peer, size_t src_off, size_t dst_off,
                                size_t bytes, int flag_index, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, 1);
    if (rc) return rc;
    CCMPI_REQUIRE(peer >= 0 && peer < lp.c[0].size, "put: peer %d out of range", peer);
    CCMPI_REQUIRE(flag_index >= 0 && flag_index + 1 < CCMPI_PAD_USER_WORDS, "put: flag %d out of range", flag_index);
    return launch(put_signal_kernel, 1, 1, (cudaStream_t)stream, lp, peer, src_off, dst_off, bytes, flag_index);
}

extern "C" int ccmpi_wait_signal(const ccmpi_dev_comm* comms, int nvr, int peer, int flag_index, void* stream) {
    Launch lp;
    int rc = fill_launch(lp, comms, 1);
    if (rc) return rc;
    CCMPI_REQUIRE(flag_index >= 0 && flag_index + 1 < CCMPI_PAD_USER_WORDS, "wait: flag %d out of range", flag_index);
    void* ptrs[] = {(void*)&lp, (void*)&flag_index};
    CCMPI_CUDA_CHECK(cudaLaunchKernel((const void*)wait_signal_kernel, dim3(1), dim3(1), ptrs, 0, (cudaStream_t)stream));
    return CCMPI_OK;
}
