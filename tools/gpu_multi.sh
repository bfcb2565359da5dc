#!/bin/bash
# Multi-GPU evidence job: run under `gpurun --gpus N -- 'bash tools/gpu_multi.sh N'`.
# Writes gpurun_out/multi_check_nN.log (DistributedCregen over NCCL == single GPU) and
# gpurun_out/bench_nN.json (the bench line at N ranks, launched the way the driver does).
N=${1:-2}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $RUN tools/multi_gpu_check.py > gpurun_out/multi_check_n$N.log 2>&1; echo "check rc=$?"; tail -5 gpurun_out/multi_check_n$N.log
timeout 600 $RUN bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; cat gpurun_out/bench_n$N.json
