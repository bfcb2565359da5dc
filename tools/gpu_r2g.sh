#!/bin/bash
# host/GPU timeline of the one-call ABI on all GPUs of the box (C3)
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1; nproc >> gpurun_out/topo_n$N.txt
timeout 300 python tools/gpu_timeline.py C3 $N > gpurun_out/timeline_n$N.txt 2>&1; cat gpurun_out/timeline_n$N.txt
