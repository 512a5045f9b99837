#!/bin/bash
# One gpurun call of the round: (1) the whole GPU suite on round 1's solver schedule (GQ_SOLVE_FUSED=0) without the fused-kernel
# tests, (2) the fused-kernel tests alone (a faulting kernel must not poison the other tests' CUDA context), (3) the suites that
# drive the solver on the default (fused) path, (4) bench.py both ways.  Everything lands in gpurun_out/<tag>_*.
tag=${1:-chk}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
run() { name=$1; shift; ( "$@" > gpurun_out/${tag}_${name}.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_${name}.log ); echo "== $name: $(tail -1 gpurun_out/${tag}_${name}.log) | $(grep -E 'passed|failed|error' gpurun_out/${tag}_${name}.log | tail -1)"; }
GQ_SOLVE_FUSED=0 run base timeout 1500 python -m pytest tests -m gpu -q -x --timeout 900 -p no:cacheprovider -k "not fused"
run fused timeout 600 python -m pytest tests/test_gpu_solver.py -m gpu -q --timeout 300 -p no:cacheprovider -k "fused"
if grep -q "rc=0" gpurun_out/${tag}_fused.log; then
  run fused_suites timeout 1200 python -m pytest tests/test_gpu_solver.py tests/test_gpu_e2e.py tests/test_gpu_quantizer.py tests/test_gpu_config_shapes.py tests/test_gpu_vs_reference_ext.py -m gpu -q --timeout 900 -p no:cacheprovider
  timeout 300 python bench.py --no-cpu > gpurun_out/${tag}_bench_fused.json 2> gpurun_out/${tag}_bench_fused.err; echo "bench fused rc=$?"
fi
GQ_SOLVE_FUSED=0 timeout 300 python bench.py --no-cpu > gpurun_out/${tag}_bench_unfused.json 2> gpurun_out/${tag}_bench_unfused.err; echo "bench unfused rc=$?"
python - <<PY
import json
for n in ('fused','unfused'):
    try:
        d=json.loads(open('gpurun_out/${tag}_bench_%s.json'%n).read().strip().splitlines()[-1])
        print(n, 'value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')}, d['clocks']['sm_mhz'])
    except Exception as e:
        print(n, 'no bench line', e)
PY
