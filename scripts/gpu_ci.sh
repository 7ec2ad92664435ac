#!/bin/bash
# GPU check run by `gpurun`: each test file in its own process under a timeout so one hung kernel
# cannot take the rest down; logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
for f in tests/test_gpu_kernels.py tests/test_gpu_gemm.py tests/test_gpu_attention.py tests/test_gpu_engine.py "$@"; do
  name=$(basename $f .py)
  timeout 600 python -m pytest $f -m gpu -q -p no:cacheprovider --timeout=300 > gpurun_out/$name.log 2>&1
  r=$?
  echo "== $f exit $r"; tail -n 25 gpurun_out/$name.log
  [ $r -ne 0 ] && rc=1
done
exit $rc
