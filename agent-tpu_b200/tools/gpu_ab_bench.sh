#!/bin/bash
# GPU box: alternate two settings of one environment variable over repeated bench.py runs on the same
# box (the GPU is power-capped, clocks differ box to box: only same-box numbers compare).
#   gpu_ab_bench.sh VAR A B [rounds=2] [extra bench args]
VAR=$1; A=$2; B=$3; ROUNDS=${4:-2}; shift 4
for r in $(seq $ROUNDS); do
  for v in "$A" "$B"; do
    env $VAR=$v python bench.py --steps 5 "$@" 2>/dev/null | python -c "
import json, sys
d = json.loads(sys.stdin.read())
k = {x['kernel']: x['ms'] for x in d['kernels']}
print('$VAR=$v', 'value', d['value'], 'e2e', d['e2e']['value'], 'MHz', d['clocks']['sm_mhz'], 'gemm', round(k['gemm_qkv'] + k['gemm_attn_out'] + k['gemm_ffn1'] + k['gemm_ffn2'], 1), 'attn', k['attention'])"
  done
done
