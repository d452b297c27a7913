"""Descriptor encodings vs the CUTLASS headers on the image (scripts/cutlass_crosscheck.cu): instruction
descriptors of the bf16 / e4m3 / block-scaled e4m3 MMAs, the scale-factor chunk layout in shared and tensor
memory, and the shared-memory descriptor tcgen05.cp reads scale chunks through. Runs on the CPU; the
compiled checker is cached under build/crosscheck and rebuilt only when its sources change (~40 s)."""
import os
import shutil
import subprocess

import pytest

from conftest import REPO

SOURCES = ["scripts/cutlass_crosscheck.cu", "scripts/cutlass_crosscheck.sh", "csrc/mx_desc.cuh"]


def _headers():
    import importlib.util
    spec = importlib.util.find_spec("flashinfer")          # located, not imported (importing it is slow)
    if spec is None or not spec.submodule_search_locations:
        return None
    d = os.path.join(list(spec.submodule_search_locations)[0], "data", "cutlass", "include")
    return d if os.path.isfile(os.path.join(d, "cute", "tensor.hpp")) else None


def test_hand_written_descriptors_match_cutlass():
    if _headers() is None or not (shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc")):
        pytest.skip("needs nvcc and the CUTLASS headers of the flashinfer wheel")
    exe = os.path.join(REPO, "build", "crosscheck", "cutlass_crosscheck")
    newest = max(os.path.getmtime(os.path.join(REPO, s)) for s in SOURCES)
    env = dict(os.environ, PATH=os.environ.get("PATH", "") + os.pathsep + "/usr/local/cuda/bin")
    if os.path.exists(exe) and os.path.getmtime(exe) >= newest:
        p = subprocess.run([exe], cwd=REPO, capture_output=True, text=True, timeout=60)
    else:
        p = subprocess.run(["bash", "scripts/cutlass_crosscheck.sh"], cwd=REPO, capture_output=True, text=True, timeout=900, env=env)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-2000:]
    assert "CROSSCHECK OK" in p.stdout and "MISMATCH" not in p.stdout
    assert p.stdout.count(" ok") >= 13
