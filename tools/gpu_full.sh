#!/bin/bash
# One-GPU evidence job: parity tests, smoke, both bench arms, launch list, one full ncu capture of the
# dominant kernel, C2/C4/C5 bench lines.  Run with: gpurun --timeout 2400 -- 'bash tools/gpu_full.sh'
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"; cat gpurun_out/bench_c3.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_tf32_kernel -c 1 -f -o gpurun_out/pair_tf32_full python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
for w in C2 C4 C5; do timeout 600 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2>gpurun_out/bench_$w.err; echo "$w rc=$?"; done
