#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tf32 or unfriendly or engine" 2>&1 | tail -4
for w in C3 C3dense C3adv C4 C4ext; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline --steps 5 > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err; echo "$w rc=$?"
  python -c "
import json; l=json.load(open('gpurun_out/bench_${w}_n1.json')); k=l['kernels']; print('$w value %.3e ms %.2f pair_ms %.2f cand %.3e frac %.2e evals %d dense %d fp64 %.1f ms | e2e %.2f pinned %.2f wall %.2f' % (l['value'], l['ms_per_step'], k['pair_stage_ms'], k['candidates'], k['candidate_fraction'], k['exact_evaluations'], k['dense_items'], k['fp64_path']['pair_stage_ms'], l['e2e']['ms_per_step'], l['e2e']['ms_per_step_pinned'], l['cregen_wall']['ms']))"
done
