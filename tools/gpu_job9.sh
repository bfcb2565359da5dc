#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gpu_tf32_check.py > gpurun_out/tf32_check.log 2>&1
echo "check rc=$?"; grep -E "same=False|violations|overflow|ALL OK|FAILED|Error|error" gpurun_out/tf32_check.log | head; tail -2 gpurun_out/tf32_check.log
for d in 0 4 5 6; do
CREST_CUDA_DEBUG=$d timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('debug=$d', 'pair_stage_ms', d['kernels']['pair_stage_ms'], 'ovf', d['kernels']['overflow'])"
done
