"""``python -m collective_communication_mpi_b200.info`` -- what this installation is and what it found on the
box, in the spirit of ``ompi_info``: versions, the native libraries and whether they load, GPUs and whether the
device plane can come up (one GPU per rank, peer access, multicast), and every ``CCMPI_*`` variable that is set.
Prints JSON with ``--json``. Touches no GPU memory and starts no communication."""
from __future__ import annotations

import json
import os
import platform
import shutil
import subprocess
import sys


def collect() -> dict:
    from . import _native as N
    from . import build as B
    out = {"package": "collective_communication_mpi_b200", "python": platform.python_version(),
           "platform": platform.platform(), "cpus": os.cpu_count()}
    try:
        import numpy
        out["numpy"] = numpy.__version__
    except Exception as e:  # noqa: BLE001
        out["numpy"] = f"unavailable: {e}"
    try:
        import torch
        out["torch"] = torch.__version__
        out["torch_cuda"] = torch.version.cuda
        ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception as e:  # noqa: BLE001
        out["torch"], ngpu = f"unavailable: {e}", 0
    libs = {}
    for name, path, loader in (("device", B.DEVICE_LIB, N.device_lib), ("host", B.HOST_LIB, N.host_lib)):
        ent = {"path": str(path), "exists": os.path.exists(path)}
        try:
            lib = loader()
            ent["loads"] = True
            if name == "device":
                ent["version"] = int(lib.ccmpi_version())
        except Exception as e:  # noqa: BLE001
            ent["loads"], ent["error"] = False, str(e)[:200]
        libs[name] = ent
    out["libraries"] = libs
    out["nvcc"] = shutil.which("nvcc") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else None)
    gpus = []
    if ngpu:
        import torch
        for i in range(ngpu):
            p = torch.cuda.get_device_properties(i)
            gpus.append({"index": i, "name": p.name, "sm": f"{p.major}.{p.minor}", "sms": p.multi_processor_count,
                         "memory_gb": round(p.total_memory / 2 ** 30, 1)})
        peers = all(torch.cuda.can_device_access_peer(0, j) for j in range(1, ngpu)) if ngpu > 1 else None
        out["peer_access_from_gpu0"] = peers
    out["gpus"] = gpus
    out["device_plane"] = ("available: one process per GPU, up to %d ranks in one NVSwitch domain" % N.MAX_RANKS) if ngpu \
        else "unavailable here (no GPU): collectives run on the shared-memory host plane, layers on CPU tensors"
    out["limits"] = {"max_ranks_device_plane": N.MAX_RANKS, "signal_pad_words": N.PAD_WORDS,
                     "flag_words_for_fused_ops": N.PAD_USER_WORDS}
    out["environment"] = {k: v for k, v in sorted(os.environ.items()) if k.startswith("CCMPI_")}
    if ngpu and shutil.which("nvidia-smi"):
        try:
            out["topology"] = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        except Exception:  # noqa: BLE001
            pass
    return out


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    info = collect()
    if "--json" in argv:
        print(json.dumps(info, indent=1))
        return 0
    topo = info.pop("topology", None)
    for k, v in info.items():
        if isinstance(v, dict):
            print(f"{k}:")
            for kk, vv in v.items():
                print(f"  {kk}: {vv}")
        elif isinstance(v, list):
            print(f"{k}: {len(v)}")
            for item in v:
                print(f"  {item}")
        else:
            print(f"{k}: {v}")
    if topo:
        print("topology (nvidia-smi topo -m):")
        print(topo)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
