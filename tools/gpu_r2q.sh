#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tf32 or golden or unfriendly or multi_pass or dense or vs_oracle or band or heavy" 2>&1 | tail -2
for w in C3 C3dense C3adv C4ext C4; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline --no-e2e --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); k=l['kernels']; print('$w default: ms %.3f pair_ms %.3f cand %.3e evals %d dense %d' % (l['ms_per_step'], k['pair_stage_ms'], k['candidates'], k['exact_evaluations'], k['dense_items']))"
done
for t in chunk=40 scanocc=2 "chunk=40,scanocc=2" scanocc=3; do
  CREST_CUDA_TUNE=$t timeout 400 python bench.py --workload C4ext --no-cpu-baseline --no-e2e --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); k=l['kernels']; print('C4ext $t: ms %.3f pair_ms %.3f cand %.3e evals %d dense %d' % (l['ms_per_step'], k['pair_stage_ms'], k['candidates'], k['exact_evaluations'], k['dense_items']))"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 20 --csv --log-file gpurun_out/launches_C4ext.csv python bench.py --workload C4ext --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(l for l in open('gpurun_out/launches_C4ext.csv') if l.startswith('"')) if r[0]!='ID']
agg=collections.OrderedDict()
for r in rows: agg.setdefault(r[4].split('(')[0][:50],[]).append(float(r[-1].replace(',','')))
for n,v in agg.items(): print(f"{n:50s} n={len(v):3d} mean={sum(v)/len(v)/1e3:9.1f} us")
PY
