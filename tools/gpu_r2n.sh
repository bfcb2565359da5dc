#!/bin/bash
# block-schedule sweep of the one-call path: e2e of C3 and C3dense per (first block %, growth gain %)
for t in default evenblocks "blockf1=15,blockgain=160" "blockf1=12,blockgain=130" "blockf1=20,blockgain=200" "blockf1=10,blockgain=150"; do for w in C3 C3dense; do
  CREST_CUDA_TUNE=$t timeout 300 python bench.py --workload $w --no-cpu-baseline --no-fp64 --steps 6 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('$w $t: e2e', round(l['e2e']['ms_per_step'],2), 'wall', round(l['cregen_wall']['ms'],2), 'step', round(l['ms_per_step'],2), 'pair', round(l['kernels']['pair_stage_ms'],2))"
done; done
