#!/bin/bash
# One 8-GPU session: multi-GPU tests, the model-parallel configurations, collective benches, the headline bench.
# Everything lands in gpurun_out/ (copied into profiles/ by hand afterwards).
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
strip() { grep -v '^\*\|OMP_NUM\|^$\|unbatched P2P' "$1"; }
bash scripts/gpu_multi_ci.sh test_symmetric_trust_allreduce_matches_oracle test_tensor_parallel_engine_matches_unsharded_engine \
  test_fused_pipeline_stages_with_reassignment test_zero1_matches_replicated_step_and_flags_poisoned_shard \
  test_k4_sum_allreduce_and_multimem_writeback test_non_finite_rank_cannot_poison_the_aggregate \
  test_trainer_zero1_flags_model_poisoning_in_the_same_step test_data_parallel_engine_replicas_stay_in_sync \
  test_trainer_fused_dp_detects_poisoned_rank
echo "=== tp collective"; timeout 300 $TR bench/tp_collective_bench.py --step > gpurun_out/tp_collective_${N}gpu.txt 2>&1; strip gpurun_out/tp_collective_${N}gpu.txt | tail -n 14 | cut -c1-260
echo "=== pp4dp2"; timeout 400 $TR bench/config_bench.py --config gpt2_pp4dp2 --steps 14 --warmup 3 --attack-step 10 > gpurun_out/config_gpt2_pp4dp2_${N}gpu.txt 2>&1; strip gpurun_out/config_gpt2_pp4dp2_${N}gpu.txt | tail -n 2 | cut -c1-1800
echo "=== tp8"; timeout 400 $TR bench/config_bench.py --config gpt2_tp8 --steps 12 --warmup 3 > gpurun_out/config_gpt2_tp8_${N}gpu.txt 2>&1; strip gpurun_out/config_gpt2_tp8_${N}gpu.txt | tail -n 2 | cut -c1-1500
echo "=== phase times"; timeout 300 $TR bench/phase_times.py > gpurun_out/phase_times_${N}gpu.txt 2>&1; strip gpurun_out/phase_times_${N}gpu.txt | head -n 34 | cut -c1-200
echo "=== bench default"; timeout 400 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; cut -c1-1200 gpurun_out/bench_${N}gpu.json; tail -n 3 gpurun_out/bench_${N}gpu.err
echo "=== bench TDL_OVERLAP=0"; TDL_OVERLAP=0 timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 --stock skip --no-overhead 2>/dev/null | tee gpurun_out/bench_${N}gpu_serial.json | cut -c1-330
echo "=== bench TDL_OVERLAP=0 TDL_BUCKETS=1"; TDL_OVERLAP=0 TDL_BUCKETS=1 timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 --stock skip --no-overhead 2>/dev/null | tee gpurun_out/bench_${N}gpu_serial_1bucket.json | cut -c1-330
