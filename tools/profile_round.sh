#!/bin/bash
# ncu evidence for one round (1 GPU), ONE ncu pass per invocation (= per gpurun call).
#   tools/profile_round.sh <tag> launches            launch list of `bench.py --steps 1 --warmup 1`
#   tools/profile_round.sh <tag> full "<target args>" <kernel regex>
#                                                    --set full capture of one kernel, e.g.  full "syrk 11008" syrk_upper
# Every ncu command runs only after the same command line exited 0 without ncu.
# Outputs land in gpurun_out/<tag>_*; summaries are copied to profiles/ by hand afterwards.
tag=${1:-prof}
mode=${2:-launches}
mkdir -p gpurun_out
if [ "$mode" = launches ]; then
  B="python bench.py --steps 1 --warmup 1 --no-cpu"
  $B > gpurun_out/${tag}_bench_plain.json 2> gpurun_out/${tag}_bench_plain.err &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
      --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_bench.log 2>&1
  echo "launch list exit $?"
else
  args=$3; pat=$4; name=$(echo $args | tr ' ' '_')
  python tools/prof_targets.py $args > gpurun_out/${tag}_${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:$pat" -s 1 -c 3 -f \
      -o gpurun_out/${tag}_${name} python tools/prof_targets.py $args > gpurun_out/${tag}_${name}_ncu.log 2>&1
  echo "$name exit $?"
fi
ls -la gpurun_out | tail -8
