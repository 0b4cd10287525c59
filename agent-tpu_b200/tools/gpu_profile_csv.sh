#!/bin/bash
# GPU box: CSV parity tests, an index-only CSV run, then the same command under ncu (full set, csv kernels).
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_csv.py -x -q -m gpu 2>&1 | tail -15
python tests/tools/gpu_csv_bench.py 1000000 0 > gpurun_out/csv_index.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'csv_|utf8_|scan_|sum_blocks' -c 12 \
    -o gpurun_out/csv_index -f python tests/tools/gpu_csv_bench.py 1000000 0 > gpurun_out/csv_ncu.log 2>&1
echo ncu_rc=$?
cat gpurun_out/csv_index.log
tail -3 gpurun_out/csv_ncu.log
