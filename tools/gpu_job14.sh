#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for t in "" "notf32"; do
CREST_CUDA_TUNE="$t" timeout 600 python bench.py --workload C4 --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('C4 tune=[$t]', '%.3e pairs/s' % d['value'], 'pair stage %.2f ms' % d['kernels']['pair_stage_ms'], d['kernels']['variant'][:28], 'cand', d['kernels']['exact_candidates'], 'ovf', d['kernels']['overflow'], 'dups', d['duplicates_found'], 'TF/s %.1f' % d['roofline']['achieved'], 'e2e %.3e' % d['e2e']['value'])"
done
