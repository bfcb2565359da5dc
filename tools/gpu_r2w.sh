#!/bin/bash
# 8 GPUs, final code: single-process engine tests, then the C3dense device step under torchrun
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "engine_single_process" > gpurun_out/engine_n8_final.log 2>&1; echo "engine rc=$?"; tail -2 gpurun_out/engine_n8_final.log
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --workload C3dense --no-fp64 --no-e2e --steps 10 > gpurun_out/bench_C3dense_n8_final.json 2> gpurun_out/bench_C3dense_n8_final.err; echo "bench rc=$?"; cat gpurun_out/bench_C3dense_n8_final.json | head -c 1500
