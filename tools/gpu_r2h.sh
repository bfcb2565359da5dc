#!/bin/bash
# e2e after the non-temporal staging copies / growing row blocks: tests of the one-call path, timeline, bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or staged or engine or full_size or vs_oracle" 2>&1 | tail -3
for t in default evenblocks copythr=4 copythr=8 copythr=12 "evenblocks,copythr=4"; do
  CREST_CUDA_TUNE=$t timeout 300 python bench.py --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('$t: e2e', round(l['e2e']['ms_per_step'],2), 'pinned', round(l['e2e']['ms_per_step_pinned'],2), 'wall', round(l['cregen_wall']['ms'],2), round(l['cregen_wall']['ms_median'],2), 'step', round(l['ms_per_step'],2))"
done
timeout 300 python tools/gpu_timeline.py C3 1 2>&1 | head -12
for w in C4 C2 C3dense; do timeout 300 python bench.py --workload $w --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('$w: e2e', round(l['e2e']['ms_per_step'],2), 'pinned', round(l['e2e']['ms_per_step_pinned'],2), 'wall', round(l['cregen_wall']['ms'],2), 'step', round(l['ms_per_step'],2))"
done
