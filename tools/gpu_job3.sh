#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gpu_tf32_check.py > gpurun_out/tf32_check.log 2>&1
echo "check rc=$?"; tail -30 gpurun_out/tf32_check.log
