tag=r02_v45
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider --durations=15 > gpurun_out/${tag}_suite.log 2>&1; echo "suite rc=$?"; tail -25 gpurun_out/${tag}_suite.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')}, d['clocks'], d['roofline']['frac'], d['gpu_launches'])
PY
