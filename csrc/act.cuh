// Activation math shared by the stand-alone elementwise kernels (elementwise.cu) and the GEMM epilogues
// that fuse them (gemm_kernel.cuh: EPI_GELU_DUAL / EPI_MUL_DGELU), so that a model gives the same numbers
// whichever path runs. tanh-approximated GELU, the form the reference-style FC block uses between fc1 and
// fc2 (the reference itself has no compute, func_impl.py:76-187).
//
// tanh comes from the SFU (`tanh.approx.f32`, relative error 2^-11): one MUFU per element instead of
// exp + reciprocal, and below the rounding step of the bf16 results these functions feed.
#pragma once

namespace act {

constexpr float kGeluC = 0.7978845608028654f;   // sqrt(2 / pi)
constexpr float kGeluA = 0.044715f;

__device__ __forceinline__ float tanh_sfu(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ float gelu(float x) {
    const float th = tanh_sfu(kGeluC * (x + kGeluA * x * x * x));
    return 0.5f * x * (1.0f + th);
}

// d gelu(x) / dx
__device__ __forceinline__ float gelu_grad(float x) {
    const float x2 = x * x;
    const float th = tanh_sfu(kGeluC * (x + kGeluA * x2 * x));
    const float dth = (1.0f - th * th) * kGeluC * (1.0f + 3.0f * kGeluA * x2);
    return 0.5f * (1.0f + th) + 0.5f * x * dth;
}

}  // namespace act
