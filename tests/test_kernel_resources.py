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
