#!/bin/bash
# final confirmation on one GPU at the last commit: whole suite, C3 bench line with the CPU baseline, the unfriendly workloads
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_C3_n1.json 2> gpurun_out/bench_C3_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_C3_n1.json | head -c 700; echo
for w in C3dense C3adv C4ext; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline --steps 5 > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err; echo "$w rc=$?"
  python -c "
import json; l=json.load(open('gpurun_out/bench_${w}_n1.json')); k=l['kernels']; print('$w value %.3e ms %.2f pair_ms %.2f cand %.3e evals %d dense %d fp64 %.1f ms | e2e %.2f pinned %.2f wall %.2f' % (l['value'], l['ms_per_step'], k['pair_stage_ms'], k['candidates'], k['exact_evaluations'], k['dense_items'], k['fp64_path']['pair_stage_ms'], l['e2e']['ms_per_step'], l['e2e']['ms_per_step_pinned'], l['cregen_wall']['ms']))"
done
