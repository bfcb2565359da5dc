#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(l for l in open('gpurun_out/launches_c.csv') if l.startswith('"')) if r[0]!='ID']
agg=collections.OrderedDict()
for r in rows: agg.setdefault(r[4].split('(')[0][:50],[]).append(float(r[-1].replace(',','')))
for n,v in agg.items(): print(f"{n:50s} n={len(v):3d} mean={sum(v)/len(v)/1e3:9.1f} us")
PY
for w in C3dense C3; do for t in default evenblocks; do
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/e2e_launches_${w}_$t.csv python tools/gpu_e2e_launches.py $w $t > /dev/null 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(l for l in open('gpurun_out/e2e_launches_${w}_$t.csv') if l.startswith('"')) if r[0]!='ID']
agg=collections.OrderedDict()
for r in rows: agg.setdefault(r[4].split('(')[0][:40],[]).append(float(r[-1].replace(',',''))/1e3)
print('== $w $t')
for n,v in agg.items(): print(f"{n:40s} n={len(v):3d} total={sum(v):9.1f} us  each: "+" ".join(f"{x:.0f}" for x in v[:8]))
PY
done; done
for w in C3 C3dense C2 C4; do timeout 300 python bench.py --workload $w --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('$w: step', round(l['ms_per_step'],3), 'pair', round(l['kernels']['pair_stage_ms'],3), 'e2e', round(l['e2e']['ms_per_step'],2), 'wall', round(l['cregen_wall']['ms'],2))"; done
