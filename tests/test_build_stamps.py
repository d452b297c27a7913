"""The native build's staleness check (collective_communication_mpi_b200/build.py). The libraries are built
in one place and loaded in another (the GPU boxes mount the repository under a different path, and eight
ranks import it at once), so a stamp must not depend on where the checkout lives -- a lesson from round 1,
when absolute include paths in the stamp made every rank rebuild and race on the first import."""
from pathlib import Path

from collective_communication_mpi_b200 import build as B


def test_stamp_ignores_include_paths_but_not_sources_flags_or_headers(tmp_path):
    src = tmp_path / "k.cu"
    src.write_text("__global__ void k() {}\n")
    a = B._stamp(src, ["-O3", "-I/root/repo/include", "-I/root/repo/csrc"], "hdr1")
    b = B._stamp(src, ["-O3", "-I/scratch/xyz/include", "-I/scratch/xyz/csrc"], "hdr1")
    assert a == b, "moving the checkout must not invalidate the objects"
    assert B._stamp(src, ["-O2", "-I/root/repo/include"], "hdr1") != a
    assert B._stamp(src, ["-O3"], "hdr2") != a
    src.write_text("__global__ void k2() {}\n")
    assert B._stamp(src, ["-O3"], "hdr1") != a


def test_in_tree_libraries_match_their_sources():
    """What travels to the GPU box must not trigger a rebuild there: after a build, nothing is stale."""
    if not B.up_to_date():          # fresh checkout: build once (every other test would do the same on import)
        B.build()
    assert B.up_to_date()
    assert Path(B.DEVICE_LIB).exists() and Path(B.HOST_LIB).exists()
