#!/bin/bash
tag=${1:-r02_v6}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 120 python tools/solve_timeline.py 4096 1024 > gpurun_out/${tag}_timeline_4096.txt 2>&1; cat gpurun_out/${tag}_timeline_4096.txt | head -48
timeout 120 python tools/solve_timeline.py 11008 1024 > gpurun_out/${tag}_timeline_11008.txt 2>&1; cat gpurun_out/${tag}_timeline_11008.txt | head -14
( timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${tag}_pytest.log | head
