#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
for w in C3 C2 C4 C3; do timeout 600 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2>gpurun_out/bench_$w.err; echo "$w rc=$?"; grep "e2e step" gpurun_out/bench_$w.err; python - <<PY
import json
d=json.load(open('gpurun_out/bench_$w.json'))
print('$w', '%.4g'%d['value'], '%.3f ms'%d['ms_per_step'], 'pair %.3f'%d['kernels']['pair_stage_ms'], 'e2e %.2f ms'%d['e2e']['ms_per_step'], 'wall %.1f'%d['cregen_wall']['ms'])
PY
done
