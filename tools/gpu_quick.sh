#!/bin/bash
# Short one-GPU job: parity tests, phase timing of the whole-CREGEN call, default bench line, launch list.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
CREST_CUDA_TIMING=1 timeout 600 python tools/gpu_wall_breakdown.py > gpurun_out/wall_breakdown.txt 2>&1; echo "wall rc=$?"; cat gpurun_out/wall_breakdown.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c3_quick.json 2> gpurun_out/bench_c3_quick.err; echo "bench rc=$?"; cat gpurun_out/bench_c3_quick.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
grep -E "pack_axis|tf32_convert|backtrack" gpurun_out/launches.csv | head -8
