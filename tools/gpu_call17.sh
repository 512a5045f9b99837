#!/bin/bash
tag=${1:-r02_v17}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/${tag}_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -2 gpurun_out/${tag}_zz_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_suite.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/${tag}_suite.log
bash tools/gpu_call16.sh ${tag}
