#!/bin/bash
# Round 2, job b: whole GPU suite + bench line with e2e through the pipelined one-call ABI.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
CREST_CUDA_TIMING=1 timeout 300 python tools/gpu_wall_breakdown.py > gpurun_out/wall_breakdown.txt 2>&1; echo "wall rc=$?"; tail -40 gpurun_out/wall_breakdown.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"; cat gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
for b in 1 2 4 8; do CREST_CUDA_TUNE=blocks=$b timeout 300 python bench.py --no-cpu-baseline --no-fp64 --steps 3 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('blocks=$b e2e', l['e2e']['ms_per_step'], 'pinned', l['e2e']['ms_per_step_pinned'], 'wall', l['cregen_wall'])"; done
