#!/bin/bash
# compute-sanitizer memcheck over the parity tests of the kernels added last (blockpair, two-phase K6,
# pack_axis with per-lane eigenproblems, staged copies); small cases only.
mkdir -p gpurun_out
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest -q -x -m gpu \
  tests/test_compare.py "tests/test_gpu_parity.py::test_group_entry" "tests/test_gpu_parity.py::test_group_nontransitive_matches" \
  "tests/test_gpu_parity.py::test_ragged_sizes" "tests/test_gpu_parity.py::test_rotconst_entry" \
  > gpurun_out/sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/sanitizer_memcheck.log | tail -3
