#!/bin/bash
# usage: tools/gpu_multi.sh <tag> <N>   (run with gpurun --gpus N)
tag=${1:-r02_multi}; N=${2:-2}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 900 $TR tools/multi_gpu_parity.py > gpurun_out/${tag}_parity_n${N}.json 2> gpurun_out/${tag}_parity_n${N}.err; echo "parity rc=$?"
tail -c 2500 gpurun_out/${tag}_parity_n${N}.json
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${tag}_bench_n${N}.json 2> gpurun_out/${tag}_bench_n${N}.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${tag}_bench_n${N}.json').read().strip().splitlines()[-1])
    print('N', d['n_gpus'], 'value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')})
    print(d['detail'].get('phase_s_by_rank'))
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/${tag}_bench_n${N}.err').read()[-2000:])
PY
