"""The allocation-free entry point ``ccmpi_gemm_rs_local_noalloc`` (csrc/gemm_rs_noalloc.cu).

It exists for hosts that cannot allocate device memory when they call into this library -- a
server that gave the whole GPU to its own arena at start-up, a plug-in living inside another
framework's memory planner, a loop that cannot take cudaMalloc's device-wide synchronisation --
so the staging buffer and flag words the GEMM -> reduce-scatter kernel needs sit in a static
device array inside the library, and the caller passes nothing but A, B and C.

This test drives it through the C ABI (ctypes on the shared library, raw pointers) and compares
with the regular entry point ``ccmpi_fused_gemm_rs`` -- same kernel, caller-provided scratch -- on
two calls in a row with different inputs, so that anything the first call leaves behind in the
shared scratch would show up in the second. The file sorts last on purpose: with ``-x`` every
other test has already run if this one fails.
"""
import ctypes as C

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF16, F32 = 1, 0


def _lib():
    from collective_communication_mpi_b200 import _native as N
    lib = N.device_lib()
    fn = lib.ccmpi_gemm_rs_local_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 9 + [C.c_void_p]
    lib.ccmpi_gemm_rs_local_noalloc_max_bytes.restype = C.c_size_t
    lib.ccmpi_gemm_rs_local_noalloc_max_bytes.argtypes = []
    return N, lib


def _noalloc(lib, a, b, out, a_km, b_kn):
    M, Nn = out.shape
    K = a.shape[0] if a_km else a.shape[1]
    return lib.ccmpi_gemm_rs_local_noalloc(
        C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(out.data_ptr()), M, Nn, K,
        a.stride(0), b.stride(0), out.stride(0), int(a_km), int(b_kn), BF16 if out.dtype == torch.bfloat16 else F32,
        C.c_void_p(torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("M,N,K,a_km,b_kn,dt", [
    (512, 512, 256, False, False, torch.bfloat16),
    (384, 256, 512, False, True, torch.float32),
    (256, 128, 320, True, True, torch.bfloat16),
])
def test_matches_the_regular_entry_point_on_two_calls_in_a_row(M, N, K, a_km, b_kn, dt):
    N_, lib = _lib()
    from collective_communication_mpi_b200.ops import fused as F
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    heap = SymmetricHeap.create_emulated(1, 64 << 20)            # scratch for the REGULAR entry point only
    comm = make_dev_comm(heap, [0], 0, 0)
    es = 2 if dt == torch.bfloat16 else 4
    slots = heap.alloc(M * N * es)
    try:
        for call in (1, 2):
            g = torch.Generator(device="cuda").manual_seed(100 * call + M)
            A = (torch.randn(M, K, device="cuda", generator=g) * 0.5).bfloat16()
            B = (torch.randn(K, N, device="cuda", generator=g) * 0.5).bfloat16()
            a = A.t().contiguous() if a_km else A
            b = B if b_kn else B.t().contiguous()
            want = torch.zeros(M, N, device="cuda", dtype=dt)
            F.gemm_rs_raw([comm], [a], [b], slots, [want], M, N, K, a_km=a_km, b_kn=b_kn, flag_base=0, epoch=call)
            got = torch.full((M, N), float("nan"), device="cuda", dtype=dt)
            rc = _noalloc(lib, a, b, got, a_km, b_kn)
            assert rc == 0, lib.ccmpi_last_error().decode()
            torch.cuda.synchronize()
            torch.testing.assert_close(got.float(), want.float(), rtol=2e-2, atol=2e-2)
            # and both agree with the plain fp32 product
            ref = A.float() @ B.float()
            torch.testing.assert_close(got.float(), ref, rtol=2e-2, atol=2e-2 * max(1.0, ref.abs().max().item()))
    finally:
        heap.close()


def test_outputs_larger_than_the_built_in_scratch_are_refused():
    N_, lib = _lib()
    cap = lib.ccmpi_gemm_rs_local_noalloc_max_bytes()
    assert cap == 16 << 20
    M, N = 4096, 4096                                              # 4096 * 4096 * 2 B = 32 MiB > 16 MiB
    a = torch.zeros(M, 64, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(N, 64, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    rc = _noalloc(lib, a, b, out, False, False)
    assert rc < 0
    msg = lib.ccmpi_last_error().decode()
    assert "built-in scratch" in msg and "ccmpi_fused_gemm_rs" in msg
