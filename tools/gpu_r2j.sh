#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for w in C3 C4 C2 C5; do for t in default packg=-2 packg=-4 packg=-8 packg=-16; do
  CREST_CUDA_TUNE=$t timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pack_axis -c 2 --csv --log-file gpurun_out/pack_${w}_$t.csv python bench.py --workload $w --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1
  echo "$w $t: $(grep pack_axis gpurun_out/pack_${w}_$t.csv | awk -F'","' '{print $5 " " $NF}' | sed 's/(.*)//' | tr '\n' ' ')"
done; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pack_axis -c 1 -f -o gpurun_out/r02_pack python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > gpurun_out/ncu_pack.log 2>&1; echo "ncu rc=$?"
timeout 300 python tools/gpu_timeline.py C3dense 1 2>&1 | grep -v "D2H dup" | head -24
timeout 600 ncu --set full --clock-control none --import-source on -k regex:cand_scan -c 1 -f -o gpurun_out/r02_scan python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > gpurun_out/ncu_scan.log 2>&1; echo "ncu scan rc=$?"
