#!/bin/bash
# GPU-box script: launch list + ncu full captures of the current kernels (round 1, second pass).
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --items 2048 --profile-run"
timeout 300 $CMD > gpurun_out/prof3_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'gemm|attention|layernorm|embed_ln|wordpiece|classify_head|ln_stats|gather|pack_offsets' -c 400 --csv --log-file gpurun_out/r01b_launches.csv $CMD > gpurun_out/ncu3_launch.log 2>&1
echo ncu_launch_exit=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm2_bf16 -s 48 -c 4 -o gpurun_out/r01b_prof_gemm2 $CMD > gpurun_out/ncu3_gemm2.log 2>&1
echo ncu_gemm2_exit=$?
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_persist -s 14 -c 1 -o gpurun_out/r01b_prof_attn $CMD > gpurun_out/ncu3_attn.log 2>&1
echo ncu_attn_exit=$?
ls -la gpurun_out | head -30
