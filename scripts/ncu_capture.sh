#!/bin/bash
# `ncu --set full` capture of every hot kernel, one instance each (B200_PROFILING.md recipe: plain run first, same command
# line, then ncu; one GPU).  Raw-page CSVs go to gpurun_out/ncu/ (copy the ones to keep into profiles/ncu/).
set -u
mkdir -p gpurun_out/ncu
CMD="python bench/ncu_targets.py"
$CMD > gpurun_out/ncu/plain.log 2>&1 || { echo "plain run failed"; tail -n 20 gpurun_out/ncu/plain.log; exit 1; }
for k in gemm_bf16_tcgen05_2cta_kernel attn_fwd_kernel attn_bwd_ps_kernel layernorm_fwd_kernel layernorm_bwd4_kernel \
         xent_cluster_kernel adamw_kernel pack_norm_kernel trust_reduce_kernel hist_kernel select_kernel moments_kernel; do
  c=1
  [ $k = gemm_bf16_tcgen05_2cta_kernel ] && c=4      # the four instantiations: plain, 2 outputs, aux-in + column sums, split-K
  [ $k = attn_bwd_ps_kernel ] && c=2                 # dQ and dK/dV kernels
  ncu --set full --clock-control none --import-source on -k regex:$k -s 0 -c $c -f -o gpurun_out/ncu/$k $CMD > gpurun_out/ncu/$k.log 2>&1
  rc=$?
  echo "$k: ncu exit $rc ($(tail -n 1 gpurun_out/ncu/$k.log | cut -c1-120))"
  if [ -f gpurun_out/ncu/$k.ncu-rep ]; then
    ncu -i gpurun_out/ncu/$k.ncu-rep --page raw --csv > gpurun_out/ncu/$k.raw.csv 2>/dev/null
    ncu -i gpurun_out/ncu/$k.ncu-rep --page details --csv > gpurun_out/ncu/$k.details.csv 2>/dev/null
    ls -la gpurun_out/ncu/$k.ncu-rep | awk '{print "  rep bytes:", $5}'
  fi
done
# keep the transfer small: only the GEMM / attention reports travel back whole (source pages); the rest as CSV
for k in layernorm_fwd_kernel layernorm_bwd4_kernel xent_cluster_kernel adamw_kernel pack_norm_kernel hist_kernel select_kernel moments_kernel; do
  rm -f gpurun_out/ncu/$k.ncu-rep
done
du -sh gpurun_out/ncu
