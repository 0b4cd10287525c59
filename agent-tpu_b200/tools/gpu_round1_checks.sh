#!/bin/bash
# GPU-box script: no-alloc test alone, full GPU suite, ncu launch list and full captures (round 1).
mkdir -p gpurun_out
set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_zz_alone.log 2>&1; echo zz_alone_exit=$?; tail -4 gpurun_out/pytest_zz_alone.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_2.log 2>&1; echo pytest_exit=$?; tail -4 gpurun_out/pytest_gpu_2.log
CMD="python bench.py --steps 1 --warmup 1 --items 2048 --profile-run"
timeout 300 $CMD > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'gemm_bf16|attention_kernel|layernorm_kernel|embed_ln|wordpiece|classify_head' -c 360 --csv --log-file gpurun_out/r01_launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo ncu1_exit=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 48 -c 4 -o gpurun_out/r01_prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1
echo ncu2_exit=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'attention_kernel|layernorm_kernel|wordpiece' -s 30 -c 3 -o gpurun_out/r01_prof_rows $CMD > gpurun_out/ncu_rows.log 2>&1
echo ncu3_exit=$?
ls -la gpurun_out/
