#!/bin/bash
# Install the UNMODIFIED reference into baseline/_ref (git-ignored; travels with gpurun).
#
# 1. The prescribed command is tried first. It fails: the reference has neither setup.py nor
#    pyproject.toml ("Directory '/root/reference' is not installable").
# 2. Fallback: copy the tree to /tmp (the mount is read-only), add a packaging-only setup.py
#    (no source file of the reference is touched), and pip-install that copy with the same
#    flags plus --no-deps (its pins -- mpi4py, h5py, pytest-mpi -- are not in the wheelhouse).
#    mpi-test.py and tests/ are not importable packages, so they are copied verbatim next to
#    the installed modules.
set -u
cd "$(dirname "$0")/.."
REF=${1:-/root/reference}
DEST=baseline/_ref
rm -rf "$DEST" && mkdir -p "$DEST"
if python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DEST" "$REF" >/tmp/ref_install_1.log 2>&1; then
  echo "prescribed install succeeded"; exit 0
fi
echo "prescribed install failed: $(tail -1 /tmp/ref_install_1.log)"
WORK=$(mktemp -d /tmp/ref_copy.XXXXXX)
cp -r "$REF"/. "$WORK"/ && chmod -R u+w "$WORK"
find "$WORK" -name __pycache__ -type d -prune -exec rm -rf {} +
cat > "$WORK/setup.py" <<'PY'
from setuptools import setup
setup(name="collective-communication-mpi-reference", version="0.0.0",
      packages=["mpi_wrapper", "model", "data"],
      package_dir={"mpi_wrapper": "mpi_wrapper", "model": "model", "data": "data"})
PY
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target "$DEST" "$WORK" >/tmp/ref_install_2.log 2>&1 \
  || { echo "fallback install failed:"; tail -5 /tmp/ref_install_2.log; exit 1; }
cp "$REF/mpi-test.py" "$DEST/mpi-test.py"
mkdir -p "$DEST/tests" && cp "$REF"/tests/*.py "$DEST/tests/"
cp "$REF/README.md" "$DEST/REFERENCE_README.md"
rm -rf "$WORK"
echo "installed into $DEST:"; ls "$DEST"
for f in mpi_wrapper/comm.py model/func_impl.py data/data_parallel_preprocess.py mpi-test.py tests/test_get_info.py; do
  cmp -s "$REF/$f" "$DEST/$f" && echo "  identical: $f" || echo "  DIFFERS: $f"
done
