#!/bin/bash
tag=${1:-r02_v10}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
( timeout 900 python -m pytest tests/test_gpu_solver.py tests/test_gpu_e2e.py tests/test_gpu_quantizer.py tests/test_gpu_config_shapes.py tests/test_native_module.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/${tag}_pytest.log | head
timeout 120 python tools/solve_timeline.py 4096 1024 > gpurun_out/${tag}_timeline_4096.txt 2>&1; cut -c1-330 gpurun_out/${tag}_timeline_4096.txt | head -24
timeout 120 python tools/solve_timeline.py 11008 1024 > gpurun_out/${tag}_timeline_11008.txt 2>&1; cut -c1-330 gpurun_out/${tag}_timeline_11008.txt | head -13
GQ_SUPERS=4 timeout 120 python tools/kernel_bench.py solve > gpurun_out/${tag}_solve_fused.jsonl 2>&1; cat gpurun_out/${tag}_solve_fused.jsonl
timeout 300 python bench.py --no-cpu --no-w4a16 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
    print('value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')}, d['clocks']['sm_mhz'], 'frac', round(d['roofline']['frac'],3))
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/${tag}_bench.err').read()[-1500:])
PY
