#!/bin/bash
# Round 2, first job: the candidate-bitmask path (tests of the TF32 path) + bench lines of the friendly and
# the unfriendly ensembles.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tf32 or dense or collapsed or multi_pass or golden or vs_oracle" > gpurun_out/pytest_tf32.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_tf32.log; tail -15 gpurun_out/pytest_tf32.log
for w in C3 C3dense C4 C4ext; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; cat gpurun_out/bench_$w.json; tail -2 gpurun_out/bench_$w.err
done
CREST_CUDA_TUNE=nodense timeout 400 python bench.py --workload C3dense --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/bench_C3dense_nodense.json 2>&1; cat gpurun_out/bench_C3dense_nodense.json
CREST_CUDA_TUNE=nodense timeout 400 python bench.py --workload C4ext --no-cpu-baseline --no-e2e --steps 5 > gpurun_out/bench_C4ext_nodense.json 2>&1; cat gpurun_out/bench_C4ext_nodense.json
