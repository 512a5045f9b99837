#!/bin/bash
# no-alloc entry point alone in a fresh process, then the whole GPU suite (-x, the no-alloc file last), then the bench
tag=${1:-r02_v15}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/${tag}_zz_alone.log 2>&1; echo "zz alone rc=$?"; tail -3 gpurun_out/${tag}_zz_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_suite.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/${tag}_suite.log
timeout 400 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/${tag}_bench.json
