#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "engine or golden or staged or unfriendly" 2>&1 | tail -2
for w in C4ext C3dense C3 C4; do
  timeout 300 python bench.py --workload $w --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('$w: e2e', round(l['e2e']['ms_per_step'],2), 'pinned', round(l['e2e']['ms_per_step_pinned'],2), 'wall', round(l['cregen_wall']['ms'],2), 'step', round(l['ms_per_step'],2))"
done
