// Descriptor encodings of the block-scaled (MXFP8) GEMM, kept in a header of plain constexpr functions so
// that scripts/cutlass_crosscheck.cu can compare them, on the host, with what the CUTLASS headers compute.
#pragma once
#include <cstdint>

namespace mx {

// Shared-memory matrix descriptor without swizzle ("version 1" = Blackwell): 8-row x 16-byte core matrices,
// `lbo` bytes apart along K and `sbo` bytes apart along M/N. The start address goes into the low 14 bits.
__host__ __device__ constexpr uint64_t smem_desc_noswizzle(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return static_cast<uint64_t>((smem_addr >> 4) & 0x3fff) | (static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fff) << 16) |
           (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fff) << 32) | (1ull << 46);
}

// Instruction descriptor of kind::mxf8f6f4.block_scale: e4m3 x e4m3, both K-major, E8M0 scales; `sf_id` is
// the byte of the tensor-memory word that holds this MMA's scale (one per 32 elements of K).
__host__ __device__ constexpr uint32_t idesc_mx(uint32_t m, uint32_t n, uint32_t sf_id) {
    return (sf_id << 4)               // B scale id
           | (0u << 7) | (0u << 10)   // A / B format: e4m3
           | ((n >> 3) << 17)
           | (1u << 23)               // scale format: E8M0
           | ((m >> 4) << 24)
           | (sf_id << 29);           // A scale id
}

// one chunk of scales for a K-block of 128 elements: 32 rows x 16 bytes, core matrices 128 bytes apart
constexpr uint32_t SF_CHUNK_LBO = 16, SF_CHUNK_SBO = 128;

}  // namespace mx
