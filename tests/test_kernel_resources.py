"""Register / stack / local-memory budget of the kernels of record, read from the built library with cuobjdump
(no GPU needed): a kernel that starts spilling or keeping arrays on the stack shows up here at build time, before
any GPU minute is spent (scripts/kernel_resources.py; table: profiles/r2_kernel_resources.txt)."""
import collections
import os
import shutil
import sys

import pytest

from conftest import REPO

sys.path.insert(0, os.path.join(REPO, "scripts"))
pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not on PATH")

# what the watchdog's trap path needs (its message arguments); anything above is spill or a stack array
TRAP_STACK = 24


@pytest.fixture(scope="module")
def families():
    import kernel_resources as K
    from collective_communication_mpi_b200 import build as B
    B.build()
    fams = collections.defaultdict(list)
    for r in K.collect():
        fams[K.family(r["name"])].append(r)
    return fams


def test_every_kernel_family_of_the_library_is_there(families):
    want = {"coll::allreduce_oneshot_kernel", "coll::allreduce_twoshot_tight_kernel", "coll::allreduce_nvls_kernel",
            "coll::allreduce_sgd_nvls_kernel", "coll::allgather_kernel", "coll::reduce_scatter_tight_kernel",
            "coll::alltoall_kernel", "coll::barrier_kernel", "coll::put_signal_kernel", "coll::wait_signal_kernel",
            "gemm::gemm_kernel_2cta<PlainPolicy>", "gemm::gemm_kernel_2cta<AgKPolicy>", "gemm::gemm_kernel_2cta<RsNPolicy>",
            "gemm::gemm_kernel_2cta<ArPolicy>", "gemm::gemm_kernel_2cta<AgOutPolicy>", "gemm::gemm_kernel_2cta<AgMPolicy>",
            "gemm::gemm_kernel_2cta<RsMPolicy>", "gemm::gemm_kernel_2cta<AgKAgOutPolicy>", "gemm::gemm_kernel_2cta<SgdPolicy>",
            "gemm::gemm_kernel_2cta<ScaledPolicy>", "mx::gemm_mxfp8_kernel", "quantize_mxfp8_kernel", "ew::gate_fwd_kernel",
            "ew::gelu_fwd_kernel", "ew::mse_kernel", "ew::sgd_sr_kernel"}
    assert want <= set(families), sorted(want - set(families))


def test_default_path_kernels_keep_nothing_on_the_stack_and_use_no_local_memory(families):
    hot = [f for f in families if f.startswith("gemm::gemm_kernel_2cta<") or f.startswith("ew::") or f in (
        "coll::allreduce_oneshot_kernel", "coll::allreduce_twoshot_tight_kernel", "coll::allreduce_nvls_kernel",
        "coll::allreduce_sgd_nvls_kernel", "coll::allgather_kernel", "coll::reduce_scatter_tight_kernel",
        "coll::alltoall_kernel", "mx::gemm_mxfp8_kernel", "quantize_mxfp8_kernel")]
    assert len(hot) >= 20
    for f in hot:
        for k in families[f]:
            assert k["stack"] <= TRAP_STACK and k["local"] == 0, (f, k["name"][:120], k["stack"], k["local"])
            assert k["reg"] <= 255


def test_reduction_blocks_fit_their_launch_bounds(families):
    # 512-thread blocks: at most 128 registers per thread fit an SM's 64 K register file
    for f in ("coll::allreduce_twoshot_tight_kernel", "coll::reduce_scatter_tight_kernel", "coll::allreduce_nvls_kernel",
              "coll::allreduce_sgd_nvls_kernel"):
        assert max(k["reg"] for k in families[f]) <= 128, f
    # the CTA-pair GEMM runs 192 threads per CTA, one CTA per SM: the full 255 registers are available to it
    assert max(k["reg"] for k in families["gemm::gemm_kernel_2cta<PlainPolicy>"]) <= 255


def test_the_generic_reductions_are_the_ones_with_stack_arrays(families):
    """Why the group-size-specialised kernels exist (and are the default): the generic ones size their in-flight
    buffers for 8 members and the compiler keeps part of them on the stack."""
    assert max(k["stack"] for k in families["coll::allreduce_twoshot_kernel"]) > 100
    assert max(k["stack"] for k in families["coll::reduce_scatter_kernel"]) > 100


# ---- SASS of the built library: the Blackwell instructions the kernels of record are made of ------------------
def _sass(mangled=None):
    import subprocess
    from collective_communication_mpi_b200 import build as B
    cmd = ["cuobjdump", "-sass"] + (["-fun", mangled] if mangled else []) + [str(B.DEVICE_LIB)]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600).stdout


KERNELS = {
    "gemm": "_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_11PlainPolicyELi2ELb0EEEvNS_6LaunchE",
    "ag_gemm": "_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_9AgKPolicyELi2ELb0EEEvNS_6LaunchE",
    "gemm_rs": "_ZN4gemm16gemm_kernel_2ctaILi256ELb0ELb1E13__nv_bfloat16NS_9RsNPolicyELi2ELb0EEEvNS_6LaunchE",
    "ag_gemm_mx": "_ZN4gemm16gemm_kernel_2ctaILi128ELb0ELb0E13__nv_bfloat16NS_9AgKPolicyELi1ELb1EEEvNS_6LaunchE",
    "allreduce_nvls": "_ZN4coll21allreduce_nvls_kernelI13__nv_bfloat16S1_EEvNS_6LaunchEmmmf",
    "allreduce_sgd_nvls": "_ZN4coll25allreduce_sgd_nvls_kernelENS_6LaunchEmmmfyjj",
}


def test_kernels_of_record_are_made_of_tcgen05_tma_and_multimem_instructions():
    import re
    found = {k: set(re.findall(r"\b(UTC[A-Z]+(?:\.[A-Z0-9_]+)*|UTMALDG(?:\.[A-Z0-9_]+)*|UBLKCP(?:\.[A-Z0-9_]+)*|LDTM(?:\.[a-zA-Z0-9_]+)*"
                               r"|LDGMC(?:\.[A-Z0-9_]+)*|[A-Z]+(?:\.[A-Z0-9_]+)*\.SYS\b)", _sass(m)))
             for k, m in KERNELS.items()}
    has = lambda k, prefix: any(x.startswith(prefix) for x in found[k])        # noqa: E731
    for k in ("gemm", "ag_gemm", "gemm_rs"):
        assert has(k, "UTCHMMA.2CTA"), (k, "tcgen05.mma cta_group::2 (bf16)")
        assert has(k, "UTMALDG.2D"), (k, "TMA tensor loads")
        assert has(k, "LDTM"), (k, "tcgen05.ld from tensor memory")
    assert has("ag_gemm", "UBLKCP") and has("gemm_rs", "UBLKCP"), "bulk copies carry the peer traffic"
    assert any(x.endswith(".SYS") for x in found["ag_gemm"]) and any(x.endswith(".SYS") for x in found["gemm_rs"]), "system-scope flags"
    assert has("ag_gemm_mx", "UTCQMMA.2CTA") and has("ag_gemm_mx", "UTCCP"), "block-scaled MMA + scale copies into tensor memory"
    assert has("allreduce_nvls", "LDGMC") and has("allreduce_sgd_nvls", "LDGMC"), "multimem.ld_reduce (reduction in the switch)"


def test_no_tma_load_in_the_library_is_a_multicast_load():
    """TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
    memory hung the GPU -- in fact the library issues no multicast TMA load at all: the only cluster-multicast
    instruction is the MMA-completion signal (tcgen05.commit -> UTCBAR), which touches shared-memory barriers."""
    import collections
    import re
    ops = collections.Counter(re.findall(r"\b((?:UTMALDG|UBLKCP|UTMASTG|UTMAREDG|UTMAPF|UTCBAR)(?:\.[A-Z0-9_]+)*)", _sass()))
    assert ops, "no SASS found"
    loads = {k: v for k, v in ops.items() if k.startswith(("UTMALDG", "UBLKCP"))}
    assert loads and not any("MULTICAST" in k or ".MC" in k for k in loads), loads
    assert any(k.startswith("UTCBAR") and "MULTICAST" in k for k in ops)
