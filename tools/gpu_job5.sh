#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gpu_tf32_check.py > gpurun_out/tf32_check.log 2>&1
echo "check rc=$?"; tail -14 gpurun_out/tf32_check.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
echo "bench rc=$?"; cat gpurun_out/bench_c3.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['kernels'], d['e2e']['value'], d['roofline']['frac'])"; tail -3 gpurun_out/bench_c3.err
