#!/bin/bash
tag=${1:-r02_v24}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 1200 python -m pytest tests/test_gpu_invert.py tests/test_gpu_solver.py tests/test_gpu_config_shapes.py tests/test_gpu_e2e.py -m gpu -x -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/${tag}_tests.log
timeout 600 python tools/invert_time.py 4096 11008 > gpurun_out/${tag}_invert_time.jsonl 2> gpurun_out/${tag}_invert_time.err; echo "invert_time rc=$?"; cat gpurun_out/${tag}_invert_time.jsonl; tail -3 gpurun_out/${tag}_invert_time.err
timeout 600 python tools/quant_time.py > gpurun_out/${tag}_quant_time.jsonl 2> gpurun_out/${tag}_quant_time.err; echo "quant_time rc=$?"; cat gpurun_out/${tag}_quant_time.jsonl
timeout 600 python tools/parity_probe.py 11008 factor > gpurun_out/${tag}_factor_sweep_11008.json 2> gpurun_out/${tag}_factor_sweep.err; echo "factor sweep rc=$?"; cat gpurun_out/${tag}_factor_sweep_11008.json
