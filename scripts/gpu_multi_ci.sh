#!/bin/bash
# Multi-GPU checks (gpurun --gpus N): every test of tests/test_gpu_multi.py in its own process under a timeout, so a
# rank that wedges cannot take the other tests down; per-test logs and a summary land in gpurun_out/.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
sum=gpurun_out/multi_${N}gpu_summary.txt
: > $sum
rc=0
tests=${@:-$(python - <<'PY'
import re
print(" ".join(re.findall(r"^def (test_\w+)", open("tests/test_gpu_multi.py").read(), re.M)))
PY
)}
for t in $tests; do
  timeout 420 python -m pytest "tests/test_gpu_multi.py::$t" -m gpu -q -x -p no:cacheprovider --timeout=400 > gpurun_out/multi_${N}gpu_$t.log 2>&1
  r=$?
  echo "$t exit $r ($(tail -n 1 gpurun_out/multi_${N}gpu_$t.log))" | tee -a $sum
  if [ $r -ne 0 ]; then rc=1; tail -n 40 gpurun_out/multi_${N}gpu_$t.log; fi
done
exit $rc
