#!/bin/bash
for d in 6 7 8; do
CREST_CUDA_DEBUG=$d timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('debug=$d', 'pair_stage_ms', d['kernels']['pair_stage_ms'], 'ovf', d['kernels']['overflow'])"
done
