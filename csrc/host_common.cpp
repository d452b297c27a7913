#include "host_common.h"

#include <vector>

#include <mutex>
#include <unordered_map>

namespace ccmpi {

static thread_local std::string g_err;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
}

void* driver_fn(const char* name) {
    static std::mutex mu;
    static std::unordered_map<std::string, void*> cache;
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(name);
    if (it != cache.end()) return it->second;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || fn == nullptr) {
        set_error("driver entry point %s unavailable (%s)", name, cudaGetErrorString(e));
        (void)cudaGetLastError();
        return nullptr;
    }
    cache[name] = fn;
    return fn;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

namespace {
struct TmapKey {
    const void* base;
    long long rows, cols, ld;
    int elem_bytes, box_rows, box_cols;
    bool operator==(const TmapKey& o) const {
        return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && elem_bytes == o.elem_bytes &&
               box_rows == o.box_rows && box_cols == o.box_cols;
    }
};
struct TmapKeyHash {
    size_t operator()(const TmapKey& k) const {
        size_t h = std::hash<const void*>()(k.base);
        auto mix = [&h](long long v) { h ^= std::hash<long long>()(v) + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2); };
        mix(k.rows); mix(k.cols); mix(k.ld); mix(k.elem_bytes); mix(k.box_rows); mix(k.box_cols);
        return h;
    }
};
}  // namespace

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, long long rows, long long cols, long long ld,
                 int box_rows, int box_cols) {
    static std::mutex mu;
    static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
    TmapKey key{base, rows, cols, ld, elem_bytes, box_rows, box_cols};
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) {
            *out = it->second;
            return CCMPI_OK;
        }
    }
    auto encode = drv<EncodeTiledFn>("cuTensorMapEncodeTiled");
    if (!encode) return CCMPI_ERR_CUDA;
    // Encoding a map over a VMM-mapped address (the symmetric heap) makes the driver look the allocation up,
    // which needs a current context. A thread that has issued no runtime call yet -- autograd's worker thread
    // entering its first backward node -- has none bound (CUDA_ERROR_INVALID_CONTEXT); bind the device's
    // primary context once per thread.
    static thread_local bool ctx_bound = false;
    if (!ctx_bound) {
        CCMPI_CUDA_CHECK(cudaFree(nullptr));
        ctx_bound = true;
    }
    CCMPI_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA operand base %p is not 16-byte aligned", base);
    CCMPI_REQUIRE((ld * elem_bytes) % 16 == 0, "TMA operand leading dimension %lld (x%d B) is not a multiple of 16 B", ld,
                  elem_bytes);
    CCMPI_REQUIRE(box_cols * elem_bytes == 128, "TMA box inner extent must be exactly one 128-byte swizzle row");
    CCMPI_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA box rows %d out of range", box_rows);
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)(ld * elem_bytes)};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8;
    CUresult r = encode(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed: CUresult %d (base %p rows %lld cols %lld ld %lld box %dx%d)", (int)r,
                  base, rows, cols, ld, box_rows, box_cols);
        return CCMPI_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 4096) cache.clear();
    cache[key] = *out;
    return CCMPI_OK;
}

int device_sm_count() {
    static int sms[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        sms[dev] = v;
    }
    return sms[dev];
}

namespace {
std::vector<int (*)(unsigned long long)>& spin_timeout_setters() {
    static std::vector<int (*)(unsigned long long)> v;     // function-local: safe during static initialisation
    return v;
}
}  // namespace
void register_spin_timeout_setter(int (*fn)(unsigned long long)) { spin_timeout_setters().push_back(fn); }

}  // namespace ccmpi

// Watchdog budget of every device-side wait, on the CURRENT device (call once per device after
// selecting it; kernels launched afterwards see the new value).
extern "C" int ccmpi_set_spin_timeout_ns(unsigned long long ns) {
    CCMPI_REQUIRE(ns >= 1000000ull, "spin timeout of %llu ns is below 1 ms", ns);
    for (auto fn : ccmpi::spin_timeout_setters()) {
        int e = fn(ns);
        if (e != 0) {
            ccmpi::set_error("cudaMemcpyToSymbol(spin timeout) failed: %s", cudaGetErrorString((cudaError_t)e));
            return CCMPI_ERR_CUDA;
        }
    }
    return CCMPI_OK;
}

extern "C" const char* ccmpi_last_error(void) { return ccmpi::g_err.c_str(); }
extern "C" int ccmpi_version(void) { return 100; }
