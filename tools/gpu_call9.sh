#!/bin/bash
tag=${1:-r02_v9}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 120 python tools/solve_timeline.py 4096 1024 > gpurun_out/${tag}_timeline_4096.txt 2>&1; cat gpurun_out/${tag}_timeline_4096.txt | cut -c1-400
timeout 300 python tools/parity_probe.py 4096 syrk > gpurun_out/${tag}_syrk_4096.json 2>&1; cat gpurun_out/${tag}_syrk_4096.json
timeout 300 python tools/parity_probe.py 11008 syrk > gpurun_out/${tag}_syrk_11008.json 2>&1; cat gpurun_out/${tag}_syrk_11008.json
( timeout 900 python -m pytest tests/test_gpu_syrk.py tests/test_gpu_config_shapes.py tests/test_gpu_vs_reference_ext.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
grep -E "^(FAILED|ERROR)|passed|failed|^E  " gpurun_out/${tag}_pytest.log | head
