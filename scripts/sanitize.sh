#!/bin/bash
# compute-sanitizer targets (the reference has no race / memory checking at all).
#   bash scripts/sanitize.sh [memcheck|racecheck|initcheck|synccheck] [driver.py]
# Default: memcheck over the zoo-op driver.  Exit code 1 on any finding; the log goes to gpurun_out/.
# (On the shared B200 pool used for this repository compute-sanitizer is switched off by the operators; there the
# out-of-bounds hunt is done with NaN guard bands around every tensor: tests/test_gpu_zoo_ops.py::*_stay_in_bounds.)
TOOL=${1:-memcheck}
DRIVER=${2:-scripts/sanitize_zoo_ops.py}
mkdir -p gpurun_out
compute-sanitizer --tool "$TOOL" --error-exitcode 1 --print-limit 20 python "$DRIVER" > "gpurun_out/sanitize_${TOOL}.log" 2>&1
rc=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize driver|Invalid|hazard" "gpurun_out/sanitize_${TOOL}.log" | head -20
exit $rc
