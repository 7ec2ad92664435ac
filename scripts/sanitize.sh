#!/bin/bash
# compute-sanitizer pass over the kernel tests (GPU box): memcheck on every test file, racecheck + synccheck on the
# hand-synchronised kernels (trust all-reduce exchange, statistics, elementwise).  Logs -> gpurun_out/sanitize_*.log
# usage: scripts/sanitize.sh [memcheck|racecheck|synccheck] [pytest args...]
tool=${1:-memcheck}; shift
mkdir -p gpurun_out
tests=${@:-tests/test_gpu_kernels.py}
timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --launch-timeout 120 \
    python -m pytest $tests -m gpu -q -x -p no:cacheprovider > gpurun_out/sanitize_$tool.log 2>&1
rc=$?
echo "== compute-sanitizer $tool on $tests: exit $rc"
grep -E "ERROR SUMMARY|passed|failed|Invalid|Race|hazard" gpurun_out/sanitize_$tool.log | tail -n 12
exit $rc
