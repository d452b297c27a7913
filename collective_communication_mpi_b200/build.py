"""In-tree native build: nvcc/g++ straight to shared libraries next to this file.

Two artefacts, both loaded with ctypes (see ``_native.py``):

* ``lib/libccmpi_b200.so`` -- device runtime + every sm_100a kernel (``csrc/*.cu``, ``csrc/*.cpp``
  except ``host_*`` transport sources). Compiled with
  ``-gencode arch=compute_100a,code=sm_100a -lineinfo`` so ncu's source page maps to our code.
  cudart is linked statically and the driver API is resolved at run time, so the library also
  loads on a machine without a GPU (the build check runs on one).
* ``lib/libccmpi_host.so`` -- CPU shared-memory transport (``csrc/host_shm.cpp``), plain g++.

The reference has no build step at all (pure Python, requirements.txt:1-12); this replaces
``pip install mpi4py`` + a system MPI.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = Path(__file__).resolve().parent / "lib"
OBJDIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-diag-suppress", "177,550",
    f"-I{INCLUDE}", f"-I{CSRC}",
]
HOST_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-Wall", "-pthread", f"-I{INCLUDE}", f"-I{CSRC}"]

DEVICE_LIB = LIBDIR / "libccmpi_b200.so"
HOST_LIB = LIBDIR / "libccmpi_host.so"


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC or put /usr/local/cuda/bin on PATH)")


def _sources():
    dev = sorted(p for p in CSRC.glob("*.cu")) + sorted(
        p for p in CSRC.glob("*.cpp") if not p.name.startswith("host_shm"))
    host = sorted(CSRC.glob("host_shm*.cpp"))
    return dev, host


def _headers_digest() -> str:
    h = hashlib.sha1()
    for p in sorted(list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list(INCLUDE.glob("*.h"))):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _stamp(src: Path, flags, hdr: str) -> str:
    """Content hash of everything that determines an object file. Include paths are left out on
    purpose: the repository is built in one place and loaded from another (the GPU boxes mount it
    under a different path), and a stamp that changed with the checkout location would trigger a
    rebuild -- by every rank at once -- on first import there."""
    h = hashlib.sha1()
    h.update(src.read_bytes())
    h.update(" ".join(f for f in flags if not f.startswith("-I")).encode())
    h.update(hdr.encode())
    return h.hexdigest()


class _BuildLock:
    """Exclusive inter-process lock: ranks that find the libraries stale build one at a time (the
    first does the work, the others see up-to-date stamps when their turn comes)."""

    def __enter__(self):
        import fcntl
        OBJDIR.mkdir(parents=True, exist_ok=True)
        self.f = open(OBJDIR / ".lock", "w")
        fcntl.flock(self.f, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl
        fcntl.flock(self.f, fcntl.LOCK_UN)
        self.f.close()


def _compile(cmd, obj: Path, stamp: str, verbose: bool):
    stamp_file = obj.with_suffix(".stamp")
    if obj.exists() and stamp_file.exists() and stamp_file.read_text() == stamp:
        return False
    if verbose:
        print("[ccmpi build]", " ".join(str(c) for c in cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compile failed: {' '.join(map(str, cmd))}\n{r.stdout}\n{r.stderr}")
    stamp_file.write_text(stamp)
    return True


def build(verbose: bool = False, force: bool = False) -> dict:
    """Compile whatever is out of date. Returns ``{"device": path, "host": path}``. Safe to call
    from several processes at once."""
    with _BuildLock():
        if not force and up_to_date():
            return {"device": DEVICE_LIB, "host": HOST_LIB}
        return _build_locked(verbose, force)


def _build_locked(verbose: bool, force: bool) -> dict:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    dev, host = _sources()
    hdr = _headers_digest()
    nvcc = _nvcc()
    gxx = shutil.which("g++") or "g++"
    jobs = []
    dev_objs, host_objs = [], []
    for src in dev:
        obj = OBJDIR / (src.name + ".o")
        dev_objs.append(obj)
        stamp = _stamp(src, NVCC_FLAGS, hdr)
        if force and obj.exists():
            obj.unlink()
        jobs.append(([nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)], obj, stamp))
    for src in host:
        obj = OBJDIR / (src.name + ".o")
        host_objs.append(obj)
        stamp = _stamp(src, HOST_FLAGS, hdr)
        if force and obj.exists():
            obj.unlink()
        jobs.append(([gxx, *HOST_FLAGS, "-c", str(src), "-o", str(obj)], obj, stamp))
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        changed = list(ex.map(lambda j: _compile(j[0], j[1], j[2], verbose), jobs))
    dev_changed = any(changed[: len(dev)]) or not DEVICE_LIB.exists()
    host_changed = any(changed[len(dev):]) or not HOST_LIB.exists()
    if dev_objs and dev_changed:
        tmp = DEVICE_LIB.with_suffix(f".so.tmp{os.getpid()}")
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(tmp), *map(str, dev_objs),
               "-cudart", "static", "-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, DEVICE_LIB)   # a process that already mapped the old file keeps a valid image
    if host_objs and host_changed:
        tmp = HOST_LIB.with_suffix(f".so.tmp{os.getpid()}")
        cmd = [gxx, "-shared", "-o", str(tmp), *map(str, host_objs), "-pthread", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, HOST_LIB)
    return {"device": DEVICE_LIB, "host": HOST_LIB}


def up_to_date() -> bool:
    dev, host = _sources()
    hdr = _headers_digest()
    for src, flags in [(s, NVCC_FLAGS) for s in dev] + [(s, HOST_FLAGS) for s in host]:
        stamp_file = (OBJDIR / (src.name + ".o")).with_suffix(".stamp")
        if not stamp_file.exists() or stamp_file.read_text() != _stamp(src, flags, hdr):
            return False
    return DEVICE_LIB.exists() and (HOST_LIB.exists() or not host)


if __name__ == "__main__":
    out = build(verbose="-v" in sys.argv, force="--force" in sys.argv)
    for k, v in out.items():
        print(f"{k}: {v}")
