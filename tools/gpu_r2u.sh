#!/bin/bash
for t in default chunk=40 blocks=1 blocks=2 "blocks=2,chunk=40" "blockf1=30,blockgain=300"; do
  CREST_CUDA_TUNE=$t timeout 300 python bench.py --workload C4ext --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('C4ext $t: e2e', round(l['e2e']['ms_per_step'],2), 'pinned', round(l['e2e']['ms_per_step_pinned'],2), 'wall', round(l['cregen_wall']['ms'],2), 'step', round(l['ms_per_step'],2))"
done
CREST_CUDA_TIMING=2 timeout 200 python tools/gpu_e2e_launches.py C4ext 2>&1 | grep -E "timeline|cregen" | head -8
