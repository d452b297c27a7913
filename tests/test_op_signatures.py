"""The layers are exercised on the CPU through ops/host_ops.py; on a GPU they call the kernel-backed
classes of ops/fused.py, ops/fused_seq.py and ops/gemm.py with the same arguments. This test pins the two
families to the same constructor / call signatures, so a layer that works against one works against the
other."""
import inspect

import pytest

torch = pytest.importorskip("torch")

from collective_communication_mpi_b200.ops import fused, fused_seq, gemm as gemm_mod, host_ops  # noqa: E402


def _params(fn, skip=("self",)):
    return [(n, p.kind, p.default) for n, p in inspect.signature(fn).parameters.items() if n not in skip]


@pytest.mark.parametrize("name,kernel_mod", [("AllGatherGemm", fused), ("GemmReduceScatter", fused), ("GemmAllReduce", fused),
                                             ("GemmAllGather", fused), ("AllGatherGemmM", fused_seq),
                                             ("GemmReduceScatterM", fused_seq)])
def test_op_classes_have_the_same_interface(name, kernel_mod):
    k, h = getattr(kernel_mod, name), getattr(host_ops, name)
    assert _params(k.__init__) == _params(h.__init__), f"{name}.__init__ differs"
    assert _params(k.__call__) == _params(h.__call__), f"{name}.__call__ differs"
    if hasattr(k, "gathered"):
        assert hasattr(h, "gathered")


@pytest.mark.parametrize("name", ["gemm", "matmul_tn", "gemm_sgd_"])
def test_gemm_functions_have_the_same_interface(name):
    assert _params(getattr(gemm_mod, name)) == _params(getattr(host_ops, name)), f"{name} differs"
