// Fused elementwise kernels of the transformer-FC block (models/fc_stack.py): the work between the
// GEMMs, each as ONE pass over HBM instead of the chain of eager ops autograd would run.
//
//   gate      a = v * sigmoid(q * k)           forward; backward recomputes sigmoid from q, k
//   gelu      tanh-approximated GELU           forward and backward
//   mse       loss = mean((y - t)^2) AND dy = 2 (y - t) / n in the same pass
//
// The reference has no compute at all (its FC layers stop at the collectives, func_impl.py:76-187);
// this is part of what the new framework adds around them. All kernels are bandwidth-bound:
// 16-byte vectors (8 bf16), fp32 math, one rounding at the end, grid = a few CTAs per SM looping
// with a grid stride.
//
// GELU forward / backward also exist as GEMM epilogues (gemm_kernel.cuh), which is what the models use;
// the kernels here serve activations that do not come straight out of a GEMM.
#include <cuda_bf16.h>

#include "act.cuh"
#include "host_common.h"
#include "sr.cuh"

namespace ew {

constexpr int THREADS = 256;

__device__ __forceinline__ uint4 ldg16(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void stg16(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }

__device__ __forceinline__ void unpack8(const uint4& r, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const float2 x = __bfloat1622float2(h[e]);
        f[2 * e] = x.x;
        f[2 * e + 1] = x.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int e = 0; e < 4; ++e) h[e] = __floats2bfloat162_rn(f[2 * e], f[2 * e + 1]);
    return r;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }

// qkv: [T][3H] (row stride ld_qkv elements), q | k | v side by side; a: [T][H] (row stride ld_a)
__global__ void __launch_bounds__(THREADS) gate_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, long long ld_qkv,
                                                           __nv_bfloat16* __restrict__ a, long long ld_a, int T, int H) {
    const int vpr = H / 8;
    const long long total = (long long)T * vpr;
    for (long long i = blockIdx.x * (long long)THREADS + threadIdx.x; i < total; i += (long long)gridDim.x * THREADS) {
        const long long t = i / vpr;
        const int c = (int)(i - t * vpr) * 8;
        const __nv_bfloat16* row = qkv + t * ld_qkv + c;
        float q[8], k[8], v[8], o[8];
        unpack8(ldg16(row), q);
        unpack8(ldg16(row + H), k);
        unpack8(ldg16(row + 2 * H), v);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = v[e] * sigmoidf_(q[e] * k[e]);
        stg16(a + t * ld_a + c, pack8(o));
    }
}

// dqkv: [T][3H] (row stride ld_d): dq | dk | dv
__global__ void __launch_bounds__(THREADS) gate_bwd_kernel(const __nv_bfloat16* __restrict__ da, long long ld_da,
                                                           const __nv_bfloat16* __restrict__ qkv, long long ld_qkv,
                                                           __nv_bfloat16* __restrict__ dqkv, long long ld_d, int T, int H) {
    const int vpr = H / 8;
    const long long total = (long long)T * vpr;
    for (long long i = blockIdx.x * (long long)THREADS + threadIdx.x; i < total; i += (long long)gridDim.x * THREADS) {
        const long long t = i / vpr;
        const int c = (int)(i - t * vpr) * 8;
        const __nv_bfloat16* row = qkv + t * ld_qkv + c;
        float q[8], k[8], v[8], g[8], dq[8], dk[8], dv[8];
        unpack8(ldg16(row), q);
        unpack8(ldg16(row + H), k);
        unpack8(ldg16(row + 2 * H), v);
        unpack8(ldg16(da + t * ld_da + c), g);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float s = sigmoidf_(q[e] * k[e]);
            const float dz = g[e] * v[e] * s * (1.0f - s);     // d loss / d (q*k)
            dq[e] = dz * k[e];
            dk[e] = dz * q[e];
            dv[e] = g[e] * s;
        }
        __nv_bfloat16* out = dqkv + t * ld_d + c;
        stg16(out, pack8(dq));
        stg16(out + H, pack8(dk));
        stg16(out + 2 * H, pack8(dv));
    }
}


__global__ void __launch_bounds__(THREADS) gelu_fwd_kernel(const __nv_bfloat16* __restrict__ u, __nv_bfloat16* __restrict__ y,
                                                           long long nvec) {
    for (long long i = blockIdx.x * (long long)THREADS + threadIdx.x; i < nvec; i += (long long)gridDim.x * THREADS) {
        float x[8], o[8];
        unpack8(ldg16(u + i * 8), x);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = act::gelu(x[e]);
        stg16(y + i * 8, pack8(o));
    }
}

__global__ void __launch_bounds__(THREADS) gelu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ u,
                                                           __nv_bfloat16* __restrict__ du, long long nvec) {
    for (long long i = blockIdx.x * (long long)THREADS + threadIdx.x; i < nvec; i += (long long)gridDim.x * THREADS) {
        float x[8], g[8], o[8];
        unpack8(ldg16(u + i * 8), x);
        unpack8(ldg16(dy + i * 8), g);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = g[e] * act::gelu_grad(x[e]);
        stg16(du + i * 8, pack8(o));
    }
}

// loss and gradient in one pass. Each CTA leaves its partial sum of squares in partial[blockIdx.x];
// the CTA that finishes last adds the partials in index order (deterministic) and writes
// loss[0] = sum / n. `ticket` must be zero on entry and is zero again on exit.
__global__ void __launch_bounds__(THREADS) mse_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ t,
                                                      __nv_bfloat16* __restrict__ dy, long long nvec, float inv_n,
                                                      float* __restrict__ partial, unsigned* __restrict__ ticket,
                                                      float* __restrict__ loss) {
    float acc = 0.0f;
    const float gs = 2.0f * inv_n;
    for (long long i = blockIdx.x * (long long)THREADS + threadIdx.x; i < nvec; i += (long long)gridDim.x * THREADS) {
        float a[8], b[8], o[8];
        unpack8(ldg16(y + i * 8), a);
        unpack8(ldg16(t + i * 8), b);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float d = a[e] - b[e];
            acc += d * d;
            o[e] = gs * d;
        }
        stg16(dy + i * 8, pack8(o));
    }
    __shared__ float warp_sum[THREADS / 32];
    __shared__ bool last;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.0f;
#pragma unroll
        for (int w = 0; w < THREADS / 32; ++w) s += warp_sum[w];
        partial[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        float s = 0.0f;
        for (unsigned b = 0; b < gridDim.x; ++b) s += *reinterpret_cast<volatile float*>(partial + b);
        loss[0] = s * inv_n;
        *ticket = 0u;
    }
}

// w -= lr * g for a flat run of bf16 weights, fp32 math, the new weight rounded stochastically (sr.cuh):
// the optimizer pass for gradients that came out of a collective. 2 reads + 1 write of 2 bytes per parameter.
// `base` = index of w[0] in the model's flat parameter numbering (random bits depend on it, not on addresses).
__global__ void __launch_bounds__(THREADS) sgd_sr_kernel(__nv_bfloat16* __restrict__ w, const __nv_bfloat16* __restrict__ g,
                                                         long long nvec, float lr, unsigned long long base, uint32_t step,
                                                         uint32_t seed) {
    for (long long i = blockIdx.x * (long long)THREADS + threadIdx.x; i < nvec; i += (long long)gridDim.x * THREADS) {
        float x[8], d[8];
        unpack8(ldg16(w + i * 8), x);
        unpack8(ldg16(g + i * 8), d);
        uint4 r;
        uint32_t* ro = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
        for (int e = 0; e < 4; ++e)
            ro[e] = sr::pack2(sr::axpy(x[2 * e], lr, d[2 * e]), sr::axpy(x[2 * e + 1], lr, d[2 * e + 1]),
                              sr::bits((base >> 1) + (unsigned long long)i * 4 + e, step, seed));
        stg16(w + i * 8, r);
    }
}

static int grid_for(long long work_items) {
    const int sms = ccmpi::device_sm_count();
    long long want = (work_items + THREADS - 1) / THREADS;
    const long long cap = (long long)sms * 8;       // 8 x 256 threads per SM: enough loads in flight for HBM
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return (int)want;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace ew

using namespace ew;

extern "C" int ccmpi_gate_fwd(const void* qkv, long long ld_qkv, void* a, long long ld_a, int T, int H, void* stream) {
    CCMPI_REQUIRE(qkv && a && T > 0 && H > 0, "gate_fwd: null pointer or empty shape");
    CCMPI_REQUIRE(H % 8 == 0 && ld_qkv % 8 == 0 && ld_a % 8 == 0 && ld_qkv >= 3LL * H && ld_a >= H && aligned16(qkv) && aligned16(a),
                  "gate_fwd: H = %d and the row strides must be multiples of 8 elements, pointers 16-byte aligned", H);
    gate_fwd_kernel<<<grid_for((long long)T * (H / 8)), THREADS, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)qkv, ld_qkv, (__nv_bfloat16*)a, ld_a, T, H);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

extern "C" int ccmpi_gate_bwd(const void* da, long long ld_da, const void* qkv, long long ld_qkv, void* dqkv,
                              long long ld_d, int T, int H, void* stream) {
    CCMPI_REQUIRE(da && qkv && dqkv && T > 0 && H > 0, "gate_bwd: null pointer or empty shape");
    CCMPI_REQUIRE(H % 8 == 0 && ld_qkv % 8 == 0 && ld_da % 8 == 0 && ld_d % 8 == 0 && ld_qkv >= 3LL * H && ld_d >= 3LL * H &&
                      ld_da >= H && aligned16(da) && aligned16(qkv) && aligned16(dqkv),
                  "gate_bwd: H = %d and the row strides must be multiples of 8 elements, pointers 16-byte aligned", H);
    gate_bwd_kernel<<<grid_for((long long)T * (H / 8)), THREADS, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)da, ld_da, (const __nv_bfloat16*)qkv, ld_qkv, (__nv_bfloat16*)dqkv, ld_d, T, H);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

extern "C" int ccmpi_gelu_fwd(const void* u, void* y, long long n, void* stream) {
    CCMPI_REQUIRE(u && y && n > 0 && n % 8 == 0 && aligned16(u) && aligned16(y),
                  "gelu_fwd: contiguous 16-byte aligned tensors with a multiple of 8 elements (n = %lld)", n);
    gelu_fwd_kernel<<<grid_for(n / 8), THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)u, (__nv_bfloat16*)y, n / 8);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

extern "C" int ccmpi_gelu_bwd(const void* dy, const void* u, void* du, long long n, void* stream) {
    CCMPI_REQUIRE(dy && u && du && n > 0 && n % 8 == 0 && aligned16(dy) && aligned16(u) && aligned16(du),
                  "gelu_bwd: contiguous 16-byte aligned tensors with a multiple of 8 elements (n = %lld)", n);
    gelu_bwd_kernel<<<grid_for(n / 8), THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)u,
                                                                           (__nv_bfloat16*)du, n / 8);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

// workspace: at least ccmpi_mse_workspace_bytes() bytes, zero-filled once by the caller before the
// first use (the kernel leaves it ready for the next call); loss: one fp32.
extern "C" size_t ccmpi_mse_workspace_bytes(void) { return 16 + sizeof(float) * 8 * 1024; }

extern "C" int ccmpi_mse_fwd_bwd(const void* y, const void* t, void* dy, long long n, void* workspace, void* loss,
                                 void* stream) {
    CCMPI_REQUIRE(y && t && dy && workspace && loss && n > 0 && n % 8 == 0 && aligned16(y) && aligned16(t) && aligned16(dy) &&
                      aligned16(workspace),
                  "mse: contiguous 16-byte aligned tensors with a multiple of 8 elements (n = %lld)", n);
    int grid = grid_for(n / 8);
    if (grid > 8 * 1024) grid = 8 * 1024;
    unsigned* ticket = reinterpret_cast<unsigned*>(workspace);
    float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + 16);
    mse_kernel<<<grid, THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)t, (__nv_bfloat16*)dy,
                                                           n / 8, 1.0f / (float)n, partial, ticket, (float*)loss);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}

// SGD step on a flat run of bf16 weights with stochastic rounding: w[i] = SR(w[i] - lr * g[i]).
extern "C" int ccmpi_sgd_sr_bf16(void* w, const void* g, long long n, float lr, unsigned long long base_index, unsigned step,
                                 unsigned seed, void* stream) {
    CCMPI_REQUIRE(w && g && n > 0 && n % 8 == 0 && aligned16(w) && aligned16(g) && base_index % 2 == 0,
                  "sgd_sr: 16-byte aligned tensors with a multiple of 8 elements (n = %lld), even base index", n);
    sgd_sr_kernel<<<grid_for(n / 8), THREADS, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)w, (const __nv_bfloat16*)g, n / 8, lr,
                                                                         base_index, step, seed);
    CCMPI_CUDA_CHECK(cudaGetLastError());
    return CCMPI_OK;
}
