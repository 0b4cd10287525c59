#!/bin/bash
# GPU-box script (one ncu invocation per call): $1 = launches | gemm2 | attn
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --items 2048 --profile-run"
case "$1" in
  launches)
    timeout 300 $CMD > gpurun_out/r01e_plain.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv \
      --log-file gpurun_out/r01e_launches.csv $CMD > gpurun_out/r01e_ncu_launch.log 2>&1 ;;
  gemm2)
    timeout 300 $CMD > gpurun_out/r01e_plain.log 2>&1 && \
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm2_bf16 -s 48 -c 4 \
      -o gpurun_out/r01e_prof_gemm2 -f $CMD > gpurun_out/r01e_ncu_gemm2.log 2>&1 ;;
  attn)
    timeout 300 $CMD > gpurun_out/r01e_plain.log 2>&1 && \
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_persist -s 14 -c 1 \
      -o gpurun_out/r01e_prof_attn -f $CMD > gpurun_out/r01e_ncu_attn.log 2>&1 ;;
esac
echo ncu_exit=$?
tail -2 gpurun_out/r01e_plain.log | cut -c1-300
