// Host-side helpers shared by the .cu / .cpp files of libccmpi_b200: error plumbing, driver
// entry points fetched through the runtime (no link-time libcuda dependency, so the library
// also loads on a box without a driver), TMA descriptor cache.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include "../include/ccmpi_b200.h"

namespace ccmpi {

void set_error(const char* fmt, ...);

#define CCMPI_CUDA_CHECK(expr)                                                                       \
    do {                                                                                             \
        cudaError_t _e = (expr);                                                                     \
        if (_e != cudaSuccess) {                                                                     \
            ::ccmpi::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return CCMPI_ERR_CUDA;                                                                   \
        }                                                                                            \
    } while (0)

#define CCMPI_DRV_CHECK(expr)                                                                        \
    do {                                                                                             \
        CUresult _r = (expr);                                                                        \
        if (_r != CUDA_SUCCESS) {                                                                    \
            ::ccmpi::set_error("%s failed: CUresult %d (%s:%d)", #expr, (int)_r, __FILE__, __LINE__); \
            return CCMPI_ERR_CUDA;                                                                   \
        }                                                                                            \
    } while (0)

#define CCMPI_REQUIRE(cond, ...)                 \
    do {                                         \
        if (!(cond)) {                           \
            ::ccmpi::set_error(__VA_ARGS__);     \
            return CCMPI_ERR_ARG;                \
        }                                        \
    } while (0)

// Resolve a driver API symbol through cudart. Returns nullptr (and sets the error) on failure.
void* driver_fn(const char* name);

template <typename Fn>
inline Fn drv(const char* name) {
    return reinterpret_cast<Fn>(driver_fn(name));
}

// A 2-D bf16/fp8 TMA descriptor for a row-major [rows][cols] tensor with leading dimension
// `ld` elements; box = [box_rows][box_cols], 128-byte swizzle. Cached by value.
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, long long rows, long long cols,
                 long long ld, int box_rows, int box_cols);

int device_sm_count();

}  // namespace ccmpi
