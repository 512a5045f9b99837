#!/bin/bash
tag=${1:-r02_v5}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
( timeout 900 python -m pytest tests/test_gpu_invert.py tests/test_gpu_solver.py tests/test_gpu_e2e.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
tail -3 gpurun_out/${tag}_pytest.log
timeout 300 python tools/parity_probe.py 4096 factor > gpurun_out/${tag}_factor_4096.json 2>&1; cat gpurun_out/${tag}_factor_4096.json
timeout 600 python tools/parity_probe.py 11008 factor > gpurun_out/${tag}_factor_11008.json 2>&1; cat gpurun_out/${tag}_factor_11008.json
GQ_SUPERS=4 timeout 120 python tools/kernel_bench.py solve invert tf32 > gpurun_out/${tag}_kernels.jsonl 2>&1; cat gpurun_out/${tag}_kernels.jsonl
timeout 600 python tools/parity_probe.py 11008 4 256 > gpurun_out/${tag}_probe_11008.json 2> gpurun_out/${tag}_probe_11008.err; cat gpurun_out/${tag}_probe_11008.json
