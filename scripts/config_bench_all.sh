#!/bin/bash
# usage: scripts/config_bench_all.sh NGPUS [extra args]   -> gpurun_out/config_<name>_<N>gpu.log
N=$1; shift
mkdir -p gpurun_out
port=29600
for c in vgg16_mp8 resnet50_dp gpt2_tp8 gpt2_pp4dp2; do
  port=$((port+1))
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench/config_bench.py --config $c "$@" > gpurun_out/config_${c}_${N}gpu.log 2>&1
  echo "== $c exit $?"; grep -E '^\{' gpurun_out/config_${c}_${N}gpu.log | cut -c1-900 || true
  grep -E "Error|error|Traceback" gpurun_out/config_${c}_${N}gpu.log | head -5
done
