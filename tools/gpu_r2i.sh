#!/bin/bash
# one-pass pack_axis: whole GPU suite, launch times per workload (new vs packold), ncu --set full of the new kernel
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for w in C3 C4 C2 C5; do for t in default packold; do
  CREST_CUDA_TUNE=$t timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pack_axis -c 3 --csv --log-file gpurun_out/pack_${w}_$t.csv python bench.py --workload $w --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1
  echo "$w $t: $(grep pack_axis gpurun_out/pack_${w}_$t.csv | awk -F'","' '{print $5 " " $NF}' | sed 's/(.*)//' | tr '\n' ' ')"
done; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pack_axis -c 1 -f -o gpurun_out/r02_pack python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --no-fp64 > gpurun_out/ncu_pack.log 2>&1; echo "ncu rc=$?"
timeout 300 python bench.py --no-cpu-baseline --no-fp64 --steps 10 > gpurun_out/bench_C3_n1_c.json 2>/dev/null; python -c "
import json; l=json.load(open('gpurun_out/bench_C3_n1_c.json')); print('C3: step', round(l['ms_per_step'],3), 'pair', round(l['kernels']['pair_stage_ms'],3), 'e2e', round(l['e2e']['ms_per_step'],2), 'wall', round(l['cregen_wall']['ms'],2))"
for t in default evenblocks; do CREST_CUDA_TUNE=$t timeout 300 python bench.py --workload C3dense --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('C3dense $t: e2e', round(l['e2e']['ms_per_step'],2), 'wall', round(l['cregen_wall']['ms'],2), 'step', round(l['ms_per_step'],2))"; done
