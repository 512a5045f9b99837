#!/bin/bash
tag=${1:-r02_v20}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 900 python -m pytest tests/test_gpu_invert.py tests/test_gpu_config_shapes.py -m gpu -x -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/${tag}_tests.log
timeout 600 python tools/invert_time.py 4096 11008 > gpurun_out/${tag}_invert_time.jsonl 2> gpurun_out/${tag}_invert_time.err; echo "invert_time rc=$?"; cat gpurun_out/${tag}_invert_time.jsonl; tail -3 gpurun_out/${tag}_invert_time.err
P="python tools/prof_targets.py"
for a in "invert 4096"; do
  n=$(echo $a | tr ' ' '_')
  $P $a > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
      --log-file gpurun_out/${tag}_launches_${n}.csv $P $a > gpurun_out/${tag}_launches_${n}.log 2>&1
  echo "launches $n exit $?"
done
