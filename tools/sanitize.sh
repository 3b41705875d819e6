#!/usr/bin/env bash
# compute-sanitizer targets for the hand-written kernels.  One tool per invocation:
#
#   tools/sanitize.sh memcheck  [pytest -k expression]
#   tools/sanitize.sh racecheck [pytest -k expression]
#   tools/sanitize.sh synccheck [pytest -k expression]
#
# Runs the GPU kernel numerics tests under the tool and leaves the report in gpurun_out/sanitize_<tool>.log.
# On the development container:  gpurun --timeout 1500 -- 'tools/sanitize.sh memcheck'
set -u
tool="${1:-memcheck}"
expr="${2:-not perf}"
mkdir -p gpurun_out
log="gpurun_out/sanitize_${tool}.log"
extra=""
if [ "$tool" = "memcheck" ]; then extra="--leak-check no --report-api-errors no"; fi
timeout ${SANITIZE_TIMEOUT:-1400} compute-sanitizer --tool "$tool" $extra --error-exitcode 99 --print-limit 20 \
    python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -x -q -m gpu -k "$expr" > "$log" 2>&1
rc=$?
tail -n 15 "$log"
echo "compute-sanitizer $tool exit code: $rc"
exit $rc
