#!/bin/bash
for w in C3adv C4ext C3dense C3 C4; do
 for t in default nodense; do
  CREST_CUDA_TUNE=$t timeout 400 python bench.py --workload $w --no-cpu-baseline --no-e2e --no-fp64 --steps 3 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); k=l['kernels']; print('$w $t pair_ms %.2f cand %.3e evals %d dense %d tried %d matched %d' % (k['pair_stage_ms'], k['candidates'], k['exact_evaluations'], k['dense_items'], k['dense_rows_tried'], k['dense_rows_matched']))"
 done
done
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "multi_pass or adversarial or unfriendly" 2>&1 | tail -4
