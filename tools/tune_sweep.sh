#!/bin/bash
# A/B sweep of the pair_tile pipeline knobs (developer tool). usage: tune_sweep.sh "<tune1>" "<tune2>" ...
for t in "$@"; do
  CREST_CUDA_TUNE="$t" timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 3 ${BENCH_ARGS} | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$t', '| %.4g pairs/s' % d['value'], '| pair_tile %.2f ms' % d['kernels']['pair_tile_ms'], '| frac %.3f' % d['roofline']['frac'], '| smem', d['kernels']['smem'], '| dups', d['duplicates_found'])"
done
