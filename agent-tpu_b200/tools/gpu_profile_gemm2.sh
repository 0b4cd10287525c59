#!/bin/bash
# GPU-box script: ncu full captures of the pair GEMM inside the pipeline, folded vs explicit LayerNorm.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --items 2048 --profile-run"
timeout 300 $CMD > gpurun_out/prof2_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm2_bf16 -s 48 -c 4 -o gpurun_out/r01_prof_gemm2_fold $CMD > gpurun_out/ncu_gemm2_fold.log 2>&1
echo ncu_fold_exit=$?
export B2_NO_LN_FOLD=1
timeout 300 $CMD > gpurun_out/prof2_plain_nofold.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm2_bf16 -s 48 -c 4 -o gpurun_out/r01_prof_gemm2_nofold $CMD > gpurun_out/ncu_gemm2_nofold.log 2>&1
echo ncu_nofold_exit=$?
ls -la gpurun_out | head -20
