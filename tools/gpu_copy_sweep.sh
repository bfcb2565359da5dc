#!/bin/bash
# Staged pageable-copy thread sweep (capi.cu big_copy): phase times of crest_cuda_cregen_cre on C3.
mkdir -p gpurun_out
for t in 0 2 4 6 8 12; do
  echo "== CREST_CUDA_COPY_THREADS=$t"
  CREST_CUDA_COPY_THREADS=$t CREST_CUDA_TIMING=1 timeout 300 python tools/gpu_wall_breakdown.py 2>&1 | grep -E "upload|D2H xyz|cregen_CRE" | tail -6
done > gpurun_out/copy_sweep.txt 2>&1
cat gpurun_out/copy_sweep.txt
