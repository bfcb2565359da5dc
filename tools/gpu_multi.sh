#!/bin/bash
# Multi-GPU evidence job: run under `gpurun --gpus N -- 'bash tools/gpu_multi.sh N [workloads...]'`.
#   gpurun_out/engine_nN.log     single-process engine tests (crest_cuda_init(N) + one-call ABI == 1 GPU)
#   gpurun_out/multi_check_nN.log  DistributedCregen over NCCL == single GPU
#   gpurun_out/bench_<W>_nN.json the bench line at N ranks, launched the way the driver does
N=${1:-2}; shift
WL=${@:-C3}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "engine" > gpurun_out/engine_n$N.log 2>&1; echo "engine rc=$?"; tail -5 gpurun_out/engine_n$N.log
timeout 300 python tools/gpu_timeline.py C3 $N > gpurun_out/timeline_n$N.txt 2>&1; tail -12 gpurun_out/timeline_n$N.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $RUN tools/multi_gpu_check.py > gpurun_out/multi_check_n$N.log 2>&1; echo "check rc=$?"; tail -5 gpurun_out/multi_check_n$N.log
for W in $WL; do
  timeout 900 $RUN bench.py --gpus $N --workload $W > gpurun_out/bench_${W}_n$N.json 2> gpurun_out/bench_${W}_n$N.err; echo "bench $W rc=$?"; cat gpurun_out/bench_${W}_n$N.json; tail -2 gpurun_out/bench_${W}_n$N.err
done
