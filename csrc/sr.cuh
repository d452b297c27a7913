// Stochastic rounding fp32 -> bf16 for optimizer updates.
//
// A bf16 weight has 8 significant bits; an SGD step lr * g is usually far below half an ulp of the weight,
// so round-to-nearest would discard it every time and training would stand still. Stochastic rounding
// rounds up with probability equal to the discarded fraction: the EXPECTED stored value is the exact fp32
// result, with no fp32 master copy and no extra memory traffic (weights stay 2 bytes, read once, written once).
//
// Random bits come from a counter-based hash of (element index, step, seed): no state, no extra loads, and
// every data-parallel replica that updates the same element in the same step draws the same bits, so
// replicas stay bit-identical.
#pragma once
#include <cstdint>

namespace sr {

__device__ __forceinline__ uint32_t hash32(uint32_t x) {     // lowbias32 (Wellons): full-avalanche 32-bit mixer
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}

// 32 random bits for element pair `pair_idx` of step `step`
__device__ __forceinline__ uint32_t bits(unsigned long long pair_idx, uint32_t step, uint32_t seed) {
    return hash32((uint32_t)pair_idx ^ hash32((uint32_t)(pair_idx >> 32) + step * 0x9E3779B9U + seed));
}

// w - lr * g with the product and the difference rounded separately (no FMA contraction), so that the
// fp32 value that gets rounded is the one a plain `w - lr * g` produces anywhere (utils/sr.py is bit-exact)
__device__ __forceinline__ float axpy(float w, float lr, float g) { return __fsub_rn(w, __fmul_rn(lr, g)); }

// fp32 -> bf16 bits with stochastic rounding; `r16` holds 16 random bits. Inf / NaN pass through truncated.
__device__ __forceinline__ uint32_t to_bf16(float x, uint32_t r16) {
    uint32_t u = __float_as_uint(x);
    if ((u & 0x7f800000U) != 0x7f800000U) u += (r16 & 0xffffU);
    return u >> 16;
}

// two values -> packed bf16x2 (lo in the low half), one 32-bit draw for the pair
__device__ __forceinline__ uint32_t pack2(float lo, float hi, uint32_t rnd) {
    return to_bf16(lo, rnd) | (to_bf16(hi, rnd >> 16) << 16);
}

}  // namespace sr
