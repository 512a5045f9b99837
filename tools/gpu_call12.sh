#!/bin/bash
tag=${1:-r02_v12}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
( timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${tag}_pytest.log | head
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
    print('value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')}, d['clocks'], 'frac', round(d['roofline']['frac'],3))
    print(d['detail'].get('rel_sq_error_by_linear'))
    print(json.dumps(d.get('w4a16'))[:1500])
    print(json.dumps(d.get('cpu_baseline'))[:900])
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/${tag}_bench.err').read()[-1500:])
PY
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/${tag}_bench_reference.json 2> gpurun_out/${tag}_bench_reference.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/${tag}_bench_reference.json
