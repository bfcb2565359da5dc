#!/bin/bash
# Round 2, final multi-GPU evidence job (run under `gpurun --gpus 8`): single-process engine tests (ngpu = 2, 4, 8 ==
# one GPU), timeline of the one-call ABI on 8 GPUs, bench lines of C3 and C3dense at 8 ranks launched the way the driver does.
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "engine" > gpurun_out/engine_n$N.log 2>&1; echo "engine rc=$?"; tail -3 gpurun_out/engine_n$N.log
timeout 200 python tools/gpu_timeline.py C3 $N 2>&1 | head -8 > gpurun_out/timeline_n$N.txt; cat gpurun_out/timeline_n$N.txt
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for W in C3 C3dense; do
  timeout 600 $RUN bench.py --gpus $N --workload $W --no-fp64 > gpurun_out/bench_${W}_n$N.json 2> gpurun_out/bench_${W}_n$N.err; echo "bench $W rc=$?"; cat gpurun_out/bench_${W}_n$N.json; grep "e2e step" gpurun_out/bench_${W}_n$N.err
done
