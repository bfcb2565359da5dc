#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cand_scan -c 1 -f -o gpurun_out/r02_scan_dense python bench.py --workload C3dense --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > gpurun_out/ncu_scan_dense.log 2>&1; echo "ncu rc=$?"
