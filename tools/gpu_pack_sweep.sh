#!/bin/bash
# pack_axis lanes-per-conformer sweep: kernel time from an ncu launch list per variant
mkdir -p gpurun_out
for g in 4 8 16; do
  CREST_CUDA_TUNE=packg=$g timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 14 --csv --log-file gpurun_out/pack_g$g.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > /dev/null 2>&1
  echo "packg=$g: $(grep pack_axis gpurun_out/pack_g$g.csv | awk -F, '{print $NF}' | tr '\n' ' ')"
done | tee gpurun_out/pack_sweep.txt
