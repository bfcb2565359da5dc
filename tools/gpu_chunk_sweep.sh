#!/bin/bash
# work-item size (j-groups per item) of pair_tf32 on one GPU: pair-stage time per variant
mkdir -p gpurun_out
for c in 256 128 64; do
  CREST_CUDA_TUNE=chunk=$c timeout 300 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('chunk=$c', 'step %.3f ms'%d['ms_per_step'], 'pair %.3f ms'%d['kernels']['pair_stage_ms'], 'items', d['kernels']['items_this_rank'])"
done | tee gpurun_out/chunk_sweep.txt
