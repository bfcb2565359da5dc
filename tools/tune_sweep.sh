#!/bin/bash
# A/B sweep of the pair kernels' knobs (CREST_CUDA_TUNE) (developer tool). usage: tune_sweep.sh "<tune1>" "<tune2>" ...
for t in "$@"; do
  CREST_CUDA_TUNE="$t" timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 10 ${BENCH_ARGS} | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$t', '| %.4g pairs/s' % d['value'], '| pair stage %.3f ms' % d['kernels']['pair_stage_ms'], '| frac %.3f' % d['roofline']['frac'], '| smem', d['kernels']['smem'], '| dups', d['duplicates_found'])"
done
