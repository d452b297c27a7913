#!/bin/bash
# Compile and run scripts/cutlass_crosscheck.cu on the CPU. Needs nvcc and the CUTLASS headers that ship
# inside the flashinfer wheel. CuTe's rotl/rotr are mutually recursive always_inline functions, which GCC
# refuses in host code; a patched copy of that one header (math.hpp) is put first on the include path.
set -e
cd "$(dirname "$0")/.."
CUT=$(python -c "import importlib.util, os; print(os.path.join(list(importlib.util.find_spec('flashinfer').submodule_search_locations)[0], 'data/cutlass/include'))")
[ -f "$CUT/cute/tensor.hpp" ] || { echo "CUTLASS headers not found under $CUT"; exit 3; }
OUT=${1:-build/crosscheck}
mkdir -p "$OUT/shim/cute/numeric"
python - "$CUT" "$OUT" <<'PY'
import sys
cut, out = sys.argv[1], sys.argv[2]
src = open(f"{cut}/cute/numeric/math.hpp").read()
a = "return static_cast<T>(s == 0 ? x : s > 0 ? (x << s) | (x >> (N - s)) : rotr(x, -s));"
b = "return static_cast<T>(s == 0 ? x : s > 0 ? (x >> s) | (x << (N - s)) : rotl(x, -s));"
assert a in src and b in src, "cute/numeric/math.hpp changed: adapt the shim"
src = src.replace(a, "return static_cast<T>(s == 0 ? x : s > 0 ? (x << s) | (x >> (N - s)) : (x >> (-s)) | (x << (N + s)));")
src = src.replace(b, "return static_cast<T>(s == 0 ? x : s > 0 ? (x >> s) | (x << (N - s)) : (x << (-s)) | (x >> (N + s)));")
open(f"{out}/shim/cute/numeric/math.hpp", "w").write(src)
PY
nvcc -std=c++17 -I"$OUT/shim" -I"$CUT" -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr \
     -o "$OUT/cutlass_crosscheck" scripts/cutlass_crosscheck.cu
"$OUT/cutlass_crosscheck" 2> "$OUT/stderr.txt"
