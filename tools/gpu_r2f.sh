#!/bin/bash
# host/GPU timeline of the one-call ABI at N=1 (C3), phase laps of crest_cuda_cregen_cre
mkdir -p gpurun_out
timeout 300 python tools/gpu_timeline.py C3 1 > gpurun_out/timeline_n1.txt 2>&1; cat gpurun_out/timeline_n1.txt
CREST_CUDA_TIMING=1 timeout 300 python tools/gpu_wall_breakdown.py > gpurun_out/wall_breakdown.txt 2>&1; tail -40 gpurun_out/wall_breakdown.txt
