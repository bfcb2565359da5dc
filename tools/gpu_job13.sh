#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py > gpurun_out/multi_check_$N.log 2>&1
echo "multi rc=$?"; grep -E "identical|MULTI|Error|error" gpurun_out/multi_check_$N.log | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err
echo "bench rc=$?"; grep '^{' gpurun_out/bench_c3_n$N.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('N=$N', d['value'], d['ms_per_step'], d['kernels']['pair_stage_ms'], d['e2e'])"
tail -3 gpurun_out/bench_c3_n$N.err
