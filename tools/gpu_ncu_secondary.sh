#!/bin/bash
# ncu --set full of the kernels around the pair stage (pack_axis, tf32_convert, exact_cand, backtrack)
# and of the f2 block-pair kernel; summaries are extracted in the container.
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'pack_axis|tf32_convert|exact_cand|backtrack' -c 4 -f -o gpurun_out/secondary_full python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_secondary.log 2>&1; echo "ncu1 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'bp_pair|bp_center|group_match' -c 3 -f -o gpurun_out/f_rows_full python tools/bench_compare.py > gpurun_out/ncu_frows.log 2>&1; echo "ncu2 rc=$?"
