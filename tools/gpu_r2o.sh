#!/bin/bash
# Round 2, final one-GPU evidence job: no-alloc test alone, whole GPU suite, smoke, both bench arms, N=1 bench lines of
# every workload, launch lists (C3, C3adv), ncu --set full of pair_tf32 / pack_axis / cand_scan, wall breakdown.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc.log 2>&1; echo "noalloc rc=$?"; tail -2 gpurun_out/pytest_noalloc.log
timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -10 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_C3_n1.json 2> gpurun_out/bench_C3_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_C3_n1.json; tail -3 gpurun_out/bench_C3_n1.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
for w in C3dense C3adv C4 C4ext C2; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline --steps 5 > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err; echo "$w rc=$?"
  python -c "
import json; l=json.load(open('gpurun_out/bench_${w}_n1.json')); k=l['kernels']; print('$w value %.3e ms %.2f pair_ms %.2f cand %.3e frac %.2e evals %d dense %d fp64 %.1f ms | e2e %.2f pinned %.2f wall %.2f' % (l['value'], l['ms_per_step'], k['pair_stage_ms'], k['candidates'], k['candidate_fraction'], k['exact_evaluations'], k['dense_items'], k['fp64_path']['pair_stage_ms'], l['e2e']['ms_per_step'], l['e2e']['ms_per_step_pinned'], l['cregen_wall']['ms']))"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fp64 > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_C3adv.csv python bench.py --workload C3adv --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1; echo "C3adv launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pair_tf32_kernel|pack_axis|cand_scan' -c 3 -f -o gpurun_out/r02_final python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
CREST_CUDA_TIMING=1 timeout 300 python tools/gpu_wall_breakdown.py > gpurun_out/wall_breakdown.txt 2>&1; tail -12 gpurun_out/wall_breakdown.txt
timeout 300 python tools/gpu_timeline.py C3 1 2>&1 | head -6 > gpurun_out/timeline_n1.txt; cat gpurun_out/timeline_n1.txt
timeout 900 python bench.py --workload C5 --no-cpu-baseline --steps 3 > gpurun_out/bench_C5_n1.json 2> gpurun_out/bench_C5_n1.err; echo "C5 rc=$?"; head -c 400 gpurun_out/bench_C5_n1.json
