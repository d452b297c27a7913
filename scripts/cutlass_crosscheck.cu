// Host-side cross-check of this repository's hand-written tcgen05 descriptor encodings against the CUTLASS /
// CuTe headers vendored on the image (flashinfer/data/cutlass/include). Nothing runs on a GPU: CuTe's
// descriptor builders are constexpr / host-callable, so their output can be compared bit for bit with
// csrc/ptx.cuh (umma_idesc) and csrc/mx_desc.cuh (block-scaled instruction descriptor, scale-chunk
// shared-memory descriptor), and the scale-factor layouts can be compared with the index formula that
// ops/gemm_fp8.py::swizzle_scales implements. Exit code 0 = everything matches.
//
//   bash scripts/cutlass_crosscheck.sh        (compiles with nvcc, runs on the CPU)
#include <cstdio>
#include <cute/tensor.hpp>
#include <cutlass/numeric_types.h>
#include <cute/arch/mma_sm100_desc.hpp>
#include <cute/arch/mma_sm100_umma.hpp>
#include <cute/atom/mma_traits_sm100.hpp>
#include <cute/atom/mma_atom.hpp>
#include <cute/atom/copy_traits_sm100.hpp>
#include <cutlass/detail/sm100_blockscaled_layout.hpp>

#include "../csrc/mx_desc.cuh"

// csrc/ptx.cuh::umma_idesc, restated (ptx.cuh is device-only apart from this function; the SASS of the
// hardware-verified bf16 kernels pins the original)
constexpr uint32_t umma_idesc(uint32_t fmt_a, uint32_t fmt_b, bool a_mn, bool b_mn, uint32_t m, uint32_t n) {
    return (1u << 4) | (fmt_a << 7) | (fmt_b << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) | ((n >> 3) << 17) |
           ((m >> 4) << 24);
}

using namespace cute;
static int failures = 0;
static void expect(const char* what, unsigned long long ours, unsigned long long theirs) {
    const bool ok = ours == theirs;
    printf("%-58s ours %016llx  cutlass %016llx  %s\n", what, ours, theirs, ok ? "ok" : "MISMATCH");
    failures += !ok;
}

int main() {
    using E4 = cutlass::float_e4m3_t;
    using SF = cutlass::float_ue8m0_t;
    using BF = cutlass::bfloat16_t;
    // ---- instruction descriptors ----
    expect("idesc kind::f16 bf16 256x256 A K-major B MN-major", umma_idesc(1, 1, false, true, 256, 256),
           uint32_t(UMMA::make_instr_desc<BF, BF, float, 256, 256, UMMA::Major::K, UMMA::Major::MN>()));
    expect("idesc kind::f16 bf16 256x128 A MN-major B MN-major", umma_idesc(1, 1, true, true, 256, 128),
           uint32_t(UMMA::make_instr_desc<BF, BF, float, 256, 128, UMMA::Major::MN, UMMA::Major::MN>()));
    expect("idesc kind::f8f6f4 e4m3 256x256 K-major", umma_idesc(0, 0, false, false, 256, 256),
           uint32_t(UMMA::make_instr_desc<E4, E4, float, 256, 256, UMMA::Major::K, UMMA::Major::K>()));
    for (uint32_t id = 0; id < 4; ++id) {
        auto d = UMMA::make_instr_desc_block_scaled<E4, E4, float, SF, 256, 128, UMMA::Major::K, UMMA::Major::K>();
        d.a_sf_id_ = id;
        d.b_sf_id_ = id;
        char name[80];
        snprintf(name, sizeof(name), "idesc kind::mxf8f6f4.block_scale 256x128, scale id %u", id);
        expect(name, mx::idesc_mx(256, 128, id), uint32_t(d));
    }
    // ---- scale-factor layouts for the 256 x 128 x 128 CTA-pair tile ----
    using MmaOp = SM100_MMA_MXF8F6F4_2x1SM_SS<E4, E4, float, SF, 256, 128, UMMA::Major::K, UMMA::Major::K>;
    auto tiled_mma = make_tiled_mma(MmaOp{});
    using TiledMma = decltype(tiled_mma);
    using Cfg = cutlass::detail::Sm1xxBlockScaledConfig<32>;
    auto sSFA = Cfg::deduce_smem_layoutSFA(tiled_mma, Shape<_256, _128, _128>{});
    auto sSFB = Cfg::deduce_smem_layoutSFB(tiled_mma, Shape<_256, _128, _128>{});
    expect("bytes of one SFA stage (this CTA's 128 rows x 4 scales)", 512, (unsigned long long)cosize(filter_zeros(sSFA)));
    expect("bytes of one SFB stage (128 columns x 4 scales)", 512, (unsigned long long)cosize(filter_zeros(sSFB)));
    // byte offset of (row r, scale k) inside a chunk: (r % 32) * 16 + (r / 32) * 4 + k
    int bad = 0;
    for (int r = 0; r < 128; ++r)
        for (int k = 0; k < 4; ++k) {
            // coordinate in CUTLASS's layout: (((r%32, r/32), 0), (any element of the 32-vector, 0)), 0, (k, 0)
            auto off = sSFA(make_coord(make_coord(make_coord(r % 32, r / 32), 0), make_coord(0, 0)), 0, make_coord(k, 0));
            bad += int(off) != (r % 32) * 16 + (r / 32) * 4 + k;
        }
    expect("scale chunk index formula, mismatching (row, k) pairs", 0, (unsigned long long)bad);
    // tensor memory: lane = r % 32 (replicated to 4 quadrants), column = r / 32, byte = k
    auto tSFA = make_tensor<typename TiledMma::FrgTypeSFA>(shape(sSFA));
    auto tl = tSFA.layout();
    bad = 0;
    for (int r = 0; r < 128; ++r)
        for (int k = 0; k < 4; ++k) {
            auto off = tl(make_coord(make_coord(make_coord(r % 32, r / 32), 0), make_coord(0, 0)), 0, make_coord(k, 0));
            // offsets are in bytes of the scale type: 4 per 32-bit column, 65536 columns per lane
            bad += (long long)off != (long long)(r % 32) * 65536 * 4 + (r / 32) * 4 + k;
        }
    expect("tensor-memory scale addressing, mismatching (row, k) pairs", 0, (unsigned long long)bad);
    // ---- shared-memory descriptor of a scale chunk as the UTCCP source ----
    {
        auto smem = make_tensor(make_smem_ptr(reinterpret_cast<SF*>(uintptr_t(0))), sSFA);
        auto smem_c = make_tensor(smem.data(), filter_zeros(smem.layout()));
        auto tmem_c = make_tensor(tSFA.data(), filter_zeros(tSFA.layout()));
        using Op = SM100::TMEM::UTCCP::SM100_UTCCP_4x32dp128bit_2cta;
        auto copy = make_utccp_copy(Op{}, tmem_c);
        auto src = copy.get_slice(0).partition_S(smem_c);
        auto desc = get_utccp_smem_desc_tensor<Op>(src);
        UMMA::SmemDescriptor d = desc(0);
        // the start address cannot be formed on the host (it is 0 here); compare everything else
        expect("UTCCP source descriptor of a scale chunk (address 0)", mx::smem_desc_noswizzle(0, mx::SF_CHUNK_LBO, mx::SF_CHUNK_SBO),
               (unsigned long long)d.desc_);
        expect("UTCCP instructions per chunk", 1, (unsigned long long)size<1>(src) * size<2>(src) * size<3>(src));
    }
    printf("%s\n", failures ? "CROSSCHECK FAILED" : "CROSSCHECK OK");
    return failures ? 1 : 0;
}
