#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer run of the C++ host transport (csrc/host_shm.cpp: shared-memory
# segment, SPSC rings, progress engine, barriers, collectives) under the real multi-process workers of the CPU
# suite ("race detection / sanitizers", the half that needs no GPU). Runs anywhere g++ has libasan.
#   bash scripts/sanitize_host.sh            -> profiles/r2_sanitize_host.txt
# The device kernels' counterpart is scripts/sanitize.sh (compute-sanitizer, needs a GPU).
set -u
cd "$(dirname "$0")/.."
OUT=${1:-profiles/r2_sanitize_host.txt}
SAN=build/san
mkdir -p "$SAN"
LIB=$PWD/$SAN/libccmpi_host_asan.so
g++ -O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined -std=c++17 -fPIC -Wall -pthread \
    -Iinclude -Icsrc -shared csrc/host_shm.cpp -o "$LIB" -lrt || exit 1
ASAN=$(gcc -print-file-name=libasan.so)
export CCMPI_HOST_LIB=$LIB LD_PRELOAD=$ASAN
# leaks: CPython never frees everything at exit; halt_on_error: a finding must fail the worker, not scroll by
export ASAN_OPTIONS=detect_leaks=0:halt_on_error=1:abort_on_error=0:exitcode=23 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1
export PYTHONPATH=$PWD CCMPI_TRANSPORT=host
{
    echo "# g++ $(g++ -dumpversion) -fsanitize=address,undefined build of csrc/host_shm.cpp; LD_PRELOAD=$ASAN"
    echo "# each line: worker, ranks, exit code, last line of its output (exit 23 = sanitizer finding)"
} > "$OUT"
run() {   # run <ranks> <worker and args...>
    local n=$1; shift
    local log=$SAN/$(echo "$*" | tr ' /' '__').log
    timeout 600 python -m collective_communication_mpi_b200.launch -n "$n" --transport host --timeout 300 -- python "$@" > "$log" 2>&1
    local rc=$?
    local findings
    findings=$(grep -c "ERROR: AddressSanitizer\|runtime error:" "$log")
    echo "[$*] ranks=$n rc=$rc sanitizer_findings=$findings  $(grep -v '^$' "$log" | tail -1 | cut -c1-160)" | tee -a "$OUT"
}
run 8 tests/workers/w_collectives.py
run 4 tests/workers/w_collectives.py --torch-cpu
run 8 tests/workers/w_nested_split.py
run 8 tests/workers/w_probe_status.py
run 8 tests/workers/w_vcollectives.py
run 4 tests/workers/w_route_agreement.py
run 4 tests/workers/w_tp_steps.py
run 4 tests/workers/w_layers.py 2 2
run 8 mpi-test.py --test_case myallreduce
run 8 mpi-test.py --test_case myalltoall
# the stress test drives the transport from pytest (many small messages, mixed communicators, ring wrap-around)
timeout 900 python -m pytest tests/test_host_stress.py -q -x > $SAN/host_stress.log 2>&1
echo "[pytest tests/test_host_stress.py] rc=$? sanitizer_findings=$(grep -c 'ERROR: AddressSanitizer\|runtime error:' $SAN/host_stress.log)  $(tail -1 $SAN/host_stress.log)" | tee -a "$OUT"
# control: the harness must SEE an error when there is one (an all-gather into a buffer that is too small, past the
# Python-side size checks)
cat > $SAN/control.py <<'PY'
import numpy as np
from mpi4py import MPI
c = MPI.COMM_WORLD
lib, h, cid, mem, n, me = c._hargs()
assert any("libccmpi_host_asan" in l for l in open("/proc/self/maps")), "sanitized library not loaded"
src = np.arange(64, dtype=np.int64)
dst = np.empty(64, dtype=np.int64)          # needs 64 * n elements
lib.ccmpi_host_allgather(h, cid, mem, n, me, src.ctypes.data, dst.ctypes.data, src.nbytes)
PY
timeout 120 python -m collective_communication_mpi_b200.launch -n 2 --transport host --timeout 60 -- python $SAN/control.py > $SAN/control.log 2>&1
echo "[control: deliberate heap overflow through ccmpi_host_allgather] rc=$? sanitizer_findings=$(grep -c 'ERROR: AddressSanitizer' $SAN/control.log)  $(grep -m1 'ERROR: AddressSanitizer' $SAN/control.log | cut -c1-120)" | tee -a "$OUT"
