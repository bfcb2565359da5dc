#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tf32 or golden or unfriendly or multi_pass or dense" 2>&1 | tail -2
for w in C3 C3dense C5; do for t in scanocc=2 scanocc=3 scanocc=4; do
  CREST_CUDA_TUNE=$t timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:cand_scan -c 24 --csv --log-file gpurun_out/scan_${w}_$t.csv python bench.py --workload $w --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1
  echo "$w $t: $(grep cand_scan gpurun_out/scan_${w}_$t.csv | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' ' | cut -c1-200)"
done; done
