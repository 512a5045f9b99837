#!/bin/bash
tag=${1:-r02_v19}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
P="python tools/prof_targets.py"
for a in "invert 4096" "invert 11008"; do
  n=$(echo $a | tr ' ' '_')
  $P $a > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
      --log-file gpurun_out/${tag}_launches_${n}.csv $P $a > gpurun_out/${tag}_launches_${n}.log 2>&1
  echo "launches $n exit $?"
done
