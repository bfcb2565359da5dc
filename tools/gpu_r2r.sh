#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in C3 C3dense C3adv C4ext C5; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --no-e2e --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); k=l['kernels']; print('$w: ms %.3f pair_ms %.3f cand %.3e evals %d dense %d tried %d matched %d' % (l['ms_per_step'], k['pair_stage_ms'], k['candidates'], k['exact_evaluations'], k['dense_items'], k['dense_rows_tried'], k['dense_rows_matched']))"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:cand_scan -c 4 --csv --log-file gpurun_out/scan_C3dense.csv python bench.py --workload C3dense --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-fp64 > /dev/null 2>&1; echo "C3dense scan: $(grep cand_scan gpurun_out/scan_C3dense.csv | awk -F'","' '{print $NF}' | tr -d '"' | tr '\n' ' ')"
timeout 300 python bench.py --workload C3dense --no-cpu-baseline --no-fp64 --steps 5 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.read()); print('C3dense e2e', round(l['e2e']['ms_per_step'],2), 'wall', round(l['cregen_wall']['ms'],2))"
