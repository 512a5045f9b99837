#!/bin/bash
tag=${1:-r02_v26}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_suite.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/${tag}_suite.log
timeout 400 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')}, d['clocks'], d['roofline']['frac'], d['gpu_launches'])
PY
B="python bench.py --steps 1 --warmup 1 --no-cpu --no-w4a16"
$B > gpurun_out/${tag}_bench_plain.json 2> gpurun_out/${tag}_bench_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_bench.log 2>&1
echo "launch list exit $?"
timeout 900 python tools/full_model.py --model 7b > gpurun_out/${tag}_full_model_7b_n1.json 2> gpurun_out/${tag}_full_model_7b_n1.err; echo "7b n1 rc=$?"; tail -c 600 gpurun_out/${tag}_full_model_7b_n1.json
