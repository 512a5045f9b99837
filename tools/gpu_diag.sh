#!/bin/bash
# Runs each GPU test file in its own process (a trapped kernel poisons only that process).
# usage: tools/gpu_diag.sh [pytest args]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv | tee gpurun_out/gpu.txt
for f in tests/test_gpu_syrk.py tests/test_gpu_invert.py tests/test_gpu_solver.py tests/test_gpu_pack_w4a16.py tests/test_gpu_e2e.py; do
  b=$(basename $f .py)
  echo "=== $f"
  timeout 600 python -m pytest $f -m gpu -q --timeout=300 "$@" > gpurun_out/$b.log 2>&1
  echo "exit $?"
  tail -25 gpurun_out/$b.log
done
