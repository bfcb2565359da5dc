#!/bin/bash
mkdir -p gpurun_out
for d in 0 4 5 6; do
CREST_CUDA_DEBUG=$d timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('debug=$d', 'pair_stage_ms', d['kernels']['pair_stage_ms'], 'ovf', d['kernels']['overflow'])"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_tf32 -c 1 -o gpurun_out/pair_tf32_full python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"
