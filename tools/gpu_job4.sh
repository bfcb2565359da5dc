#!/bin/bash
mkdir -p gpurun_out
python - > gpurun_out/h2d.log 2>&1 <<'PY'
import torch, time, numpy as np
x = torch.empty(144_000_000 // 8, dtype=torch.float64).pin_memory()
d = torch.empty_like(x, device="cuda")
for _ in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(x); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("pinned H2D 144MB: %.2f ms  %.1f GB/s" % (dt * 1e3, 0.144 / dt))
y = torch.empty(144_000_000 // 8, dtype=torch.float64)
for _ in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(y); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("pageable H2D 144MB: %.2f ms  %.1f GB/s" % (dt * 1e3, 0.144 / dt))
PY
cat gpurun_out/h2d.log
timeout 600 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
echo "bench rc=$?"; cat gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_tf32 -c 1 -o gpurun_out/pair_tf32_full python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"
