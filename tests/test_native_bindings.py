"""The ctypes declarations load against the built libraries without a GPU: every symbol named in
``_native.py`` exists in ``libccmpi_b200.so`` / ``libccmpi_host.so`` (a typo or a stale build fails here,
on the CPU, instead of on the first GPU box)."""
from collective_communication_mpi_b200 import _native as N


def test_device_and_host_libraries_load_and_every_declared_symbol_resolves():
    dev, host = N.device_lib(), N.host_lib()
    assert dev.ccmpi_version() >= 100
    for name in ("ccmpi_gemm_bf16_ex", "ccmpi_fused_ag_gemm_ex", "ccmpi_fused_gemm_rs_ex", "ccmpi_fused_ag_ctas",
                 "ccmpi_set_spin_timeout_ns", "ccmpi_allreduce"):
        assert getattr(dev, name).argtypes is not None, name
    assert host.ccmpi_host_reserve_comm_ids.argtypes is not None
