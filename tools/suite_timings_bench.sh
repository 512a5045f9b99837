#!/bin/bash
tag=${1:-r02_v28}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_suite.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/${tag}_suite.log
timeout 600 python tools/quant_time.py > gpurun_out/${tag}_quant_time.jsonl 2> gpurun_out/${tag}_quant_time.err; echo "quant_time rc=$?"; cat gpurun_out/${tag}_quant_time.jsonl; tail -2 gpurun_out/${tag}_quant_time.err
timeout 600 python tools/invert_time.py 4096 11008 > gpurun_out/${tag}_invert_time.jsonl 2> gpurun_out/${tag}_invert_time.err; cat gpurun_out/${tag}_invert_time.jsonl
timeout 400 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')}, d['clocks'], d['roofline']['frac'], d['gpu_launches'])
PY
