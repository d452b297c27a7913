"""ctypes bindings for the two native libraries (see include/ccmpi_b200.h, include/ccmpi_host.h).

Loading policy: use the in-tree ``lib/*.so`` when it matches the sources, otherwise build it
(nvcc is on every box this runs on). A missing or unloadable device library on a machine that
has a GPU is a hard error -- there is no silent PyTorch fallback for the device ops.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path

from . import build as _build

MAX_RANKS = 8
MAX_VRANKS = 4
MAX_BLOCKS = 160
BLOCKS_PAIRED = 0x10000     # OR into max_blocks: even grid launched as 2-CTA clusters (whole SM pairs)
PAD_WORDS = 16384
PAD_BYTES = PAD_WORDS * 4
PAD_USER = 2048
PAD_USER_WORDS = PAD_WORDS - PAD_USER

# ccmpi_dtype
F32, BF16, F16, F64, I32, I64, U8 = range(7)
DTYPE_SIZE = {F32: 4, BF16: 2, F16: 2, F64: 8, I32: 4, I64: 8, U8: 1}
# ccmpi_op
SUM, MIN, MAX = range(3)
# ccmpi_ar_algo
AR_AUTO, AR_ONESHOT, AR_TWOSHOT, AR_NVLS, AR_TWOSHOT_GENERIC = range(5)

ERR_UNSUPPORTED = -3


class DevComm(C.Structure):
    """Mirror of ``ccmpi_dev_comm``."""
    _fields_ = [
        ("rank", C.c_int32),
        ("size", C.c_int32),
        ("heap", C.c_void_p * MAX_RANKS),
        ("mc_heap", C.c_void_p),
        ("pad", C.c_void_p * MAX_RANKS),
    ]


class NativeError(RuntimeError):
    pass


# Number of kernels of THIS library launched by this process (every python wrapper that ends in a
# kernel launch bumps it); bench.py reports the count inside its timed region.
LAUNCHES = [0]


def count_launch(n: int = 1) -> None:
    LAUNCHES[0] += n


_lock = threading.Lock()
_dev = None
_host = None


def _ensure_built():
    if os.environ.get("CCMPI_NO_BUILD") == "1":
        return
    if not _build.up_to_date():
        _build.build(verbose=os.environ.get("CCMPI_BUILD_VERBOSE") == "1")


def device_lib():
    """The device library (runtime + kernels), loaded once per process."""
    global _dev
    if _dev is not None:
        return _dev
    with _lock:
        if _dev is not None:
            return _dev
        _ensure_built()
        path = _build.DEVICE_LIB
        if not Path(path).exists():
            raise NativeError(f"{path} is missing; run `python -m collective_communication_mpi_b200.build`")
        lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
        _declare_device(lib)
        _dev = lib
        return lib


def host_lib():
    global _host
    if _host is not None:
        return _host
    with _lock:
        if _host is not None:
            return _host
        _ensure_built()
        # CCMPI_HOST_LIB: another build of the same sources, e.g. the AddressSanitizer one of scripts/sanitize_host.sh
        path = os.environ.get("CCMPI_HOST_LIB") or _build.HOST_LIB
        if not Path(path).exists():
            raise NativeError(f"{path} is missing; run `python -m collective_communication_mpi_b200.build`")
        lib = C.CDLL(str(path))
        _declare_host(lib)
        _host = lib
        return lib


_timeout_done = set()


def apply_device_timeout(device: int, seconds: float = None) -> float:
    """Set the watchdog budget of the device-side waits on ``device`` (once per device and process
    unless ``seconds`` is given): ``CCMPI_DEVICE_TIMEOUT_S``, default 120 s -- the same default as the
    host plane's ``CCMPI_TIMEOUT``. A rank may reach a collective long after its peers (data loading,
    first-call module load, a debugger); the peers' kernels wait that long before they trap."""
    if seconds is None:
        if device in _timeout_done:
            return -1.0
        seconds = float(os.environ.get("CCMPI_DEVICE_TIMEOUT_S", "120"))
    import torch
    with torch.cuda.device(device):
        check(device_lib().ccmpi_set_spin_timeout_ns(int(seconds * 1e9)), "ccmpi_set_spin_timeout_ns")
    _timeout_done.add(device)
    return seconds


def check(rc: int, what: str = "") -> int:
    """Raise on a negative native status, with the library's own message."""
    if rc is not None and rc < 0:
        msg = device_lib().ccmpi_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed ({rc}): {msg}")
    return rc


def check_host(rc: int, what: str = "") -> int:
    if rc is not None and rc < 0:
        msg = host_lib().ccmpi_host_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'host transport call'} failed ({rc}): {msg}")
    return rc


def _declare_device(lib):
    vp, i, i64, sz, f = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
    P = C.POINTER
    PV = P(vp)
    u = C.c_uint
    sigs = {
        "ccmpi_last_error": (C.c_char_p, []),
        "ccmpi_version": (i, []),
        "ccmpi_set_spin_timeout_ns": (i, [C.c_ulonglong]),
        "ccmpi_probe": (i, [i, P(C.c_int)]),
        "ccmpi_heap_create": (i64, [i, sz]),
        "ccmpi_heap_create_emulated": (i64, [i, sz, i]),
        "ccmpi_heap_bytes": (sz, [i64]),
        "ccmpi_heap_local_ptr": (vp, [i64]),
        "ccmpi_heap_listen": (i, [i64, C.c_char_p, i]),
        "ccmpi_heap_connect_all": (i, [i64, C.c_char_p, i, i]),
        "ccmpi_heap_peer_ptr": (vp, [i64, i]),
        "ccmpi_mc_create": (i, [i64, C.c_char_p, i, P(C.c_int), i, i]),
        "ccmpi_mc_bind": (i, [i64, i]),
        "ccmpi_mc_ptr": (vp, [i64, i]),
        "ccmpi_heap_destroy": (i, [i64]),
        "ccmpi_barrier": (i, [P(DevComm), i, vp]),
        "ccmpi_allreduce": (i, [P(DevComm), i, i, i, i, sz, sz, sz, sz, f, i, i, vp]),
        "ccmpi_allgather": (i, [P(DevComm), i, sz, sz, sz, sz, sz, sz, i, i, vp]),
        "ccmpi_reduce_scatter": (i, [P(DevComm), i, i, i, sz, sz, sz, sz, sz, sz, i, i, vp]),
        "ccmpi_alltoall": (i, [P(DevComm), i, sz, sz, sz, i, i, vp]),
        "ccmpi_put_signal": (i, [P(DevComm), i, i, sz, sz, sz, i, vp]),
        "ccmpi_wait_signal": (i, [P(DevComm), i, i, i, vp]),
        "ccmpi_gemm_bf16": (i, [vp, vp, vp, i, i, i, i, i, i, i, i, i, i, i, vp]),
        "ccmpi_gemm_bf16_ex": (i, [vp, vp, vp, i, i, i, i, i, i, i, i, i, i, i, vp, vp, i64, vp, i, vp]),
        "ccmpi_fused_ag_gemm_ex": (i, [P(DevComm), i, sz, i, sz, PV, PV, i, i, i, i, i, i, i, i, u, u, i, PV, PV, i64, vp, i,
                                       i, i, vp]),
        "ccmpi_fused_gemm_rs_ex": (i, [P(DevComm), i, PV, PV, sz, PV, i, i, i, i, i, i, i, i, i, i, u, PV, i64, vp, i, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    u = C.c_uint
    fused = {
        "ccmpi_fused_ag_gemm": (i, [P(DevComm), i, sz, i, sz, PV, PV, i, i, i, i, i, i, i, i, u, u, i, i, i, vp]),
        "ccmpi_fused_ag_ctas": (i, [i, i, i, i, i]),
        "ccmpi_fused_gemm_rs": (i, [P(DevComm), i, PV, PV, sz, PV, i, i, i, i, i, i, i, i, i, i, u, i, vp]),
        "ccmpi_fused_gemm_ar": (i, [P(DevComm), i, PV, PV, sz, sz, i, i, i, i, i, i, i, i, i, u, i, i, vp]),
        "ccmpi_fused_gemm_ag": (i, [P(DevComm), i, PV, PV, sz, i, i, i, i, i, i, i, i, i, i, vp]),
        "ccmpi_fused_ag_gemm_m": (i, [P(DevComm), i, sz, i, sz, PV, PV, i, i, i, i, i, i, i, i, u, u, i, vp]),
        "ccmpi_fused_gemm_rs_m": (i, [P(DevComm), i, PV, PV, sz, PV, i, i, i, i, i, i, i, i, i, i, u, i, vp]),
        "ccmpi_fused_ag_gemm_ag": (i, [P(DevComm), i, sz, i, sz, PV, sz, i, i, i, i, i, i, i, u, u, i, vp]),
    }
    ll = C.c_longlong
    fused.update({
        "ccmpi_gate_fwd": (i, [vp, ll, vp, ll, i, i, vp]),
        "ccmpi_gate_bwd": (i, [vp, ll, vp, ll, vp, ll, i, i, vp]),
        "ccmpi_gelu_fwd": (i, [vp, vp, ll, vp]),
        "ccmpi_gelu_bwd": (i, [vp, vp, vp, ll, vp]),
        "ccmpi_reduce_scatter_tight": (i, [P(DevComm), i, i, i, sz, sz, sz, sz, sz, sz, i, vp]),
        "ccmpi_allreduce_twoshot_tight": (i, [P(DevComm), i, i, i, sz, sz, sz, sz, f, i, vp]),
        "ccmpi_gemm_sgd_bf16": (i, [vp, vp, vp, i, i, i, i, i, i, i, i, f, i, vp]),
        "ccmpi_gemm_mx": (i, [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i, vp]),
        "ccmpi_fused_ag_gemm_mx": (i, [P(DevComm), i, sz, i, sz, sz, sz, PV, PV, PV, i, i, i, i, i, i, i, u, u, vp, i, vp]),
        "ccmpi_fused_gemm_rs_mx": (i, [P(DevComm), i, PV, PV, PV, PV, sz, PV, i, i, i, i, i, i, i, i, u, PV, i64, vp, i, vp]),
        "ccmpi_quantize_mxfp8": (i, [vp, C.c_longlong, vp, C.c_longlong, vp, i, i, vp]),
        "ccmpi_gemm_sgd_bf16_sr": (i, [vp, vp, vp, i, i, i, i, i, i, i, i, f, i, u, i, vp]),
        "ccmpi_sgd_sr_bf16": (i, [vp, vp, C.c_longlong, f, C.c_ulonglong, u, u, vp]),
        "ccmpi_allreduce_sgd_nvls": (i, [P(DevComm), i, sz, sz, sz, f, C.c_ulonglong, u, u, i, vp]),
        "ccmpi_gemm_mxfp8": (i, [vp, vp, vp, vp, vp, i, i, i, i, i, i, i, i, vp]),
        "ccmpi_mse_workspace_bytes": (sz, []),
        "ccmpi_mse_fwd_bwd": (i, [vp, vp, vp, ll, vp, vp, vp]),
    })
    for name, (res, args) in fused.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def _declare_host(lib):
    vp, i, i64, sz, d = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_double
    P = C.POINTER
    sigs = {
        "ccmpi_host_last_error": (C.c_char_p, []),
        "ccmpi_host_attach": (i64, [C.c_char_p, i, i, sz, sz, d]),
        "ccmpi_host_detach": (i, [i64, i]),
        "ccmpi_host_barrier": (i, [i64, i, P(C.c_int), i, i]),
        "ccmpi_host_allreduce": (i, [i64, i, P(C.c_int), i, i, vp, vp, sz, i, i]),
        "ccmpi_host_allgather": (i, [i64, i, P(C.c_int), i, i, vp, vp, sz]),
        "ccmpi_host_reduce_scatter": (i, [i64, i, P(C.c_int), i, i, vp, vp, sz, i, i]),
        "ccmpi_host_alltoall": (i, [i64, i, P(C.c_int), i, i, vp, vp, sz]),
        "ccmpi_host_send": (i, [i64, i, i, i, vp, sz]),
        "ccmpi_host_recv": (i, [i64, i, i, i, vp, sz]),
        "ccmpi_host_isend": (i64, [i64, i, i, i, vp, sz]),
        "ccmpi_host_irecv": (i64, [i64, i, i, i, vp, sz]),
        "ccmpi_host_test": (i, [i64, i64]),
        "ccmpi_host_waitall": (i, [i64, P(C.c_int64), i]),
        "ccmpi_host_probe": (i64, [i64, i]),
        "ccmpi_host_iprobe": (i, [i64, i, P(C.c_int), P(C.c_int), P(C.c_int64)]),
        "ccmpi_host_reserve_comm_ids": (i, [i64, i]),
        "ccmpi_host_ack_post": (i, [i64, i]),
        "ccmpi_host_ack_count": (i64, [i64, i]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
