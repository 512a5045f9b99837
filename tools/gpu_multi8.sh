#!/bin/bash
# usage: tools/gpu_multi8.sh <tag>   (run with gpurun --gpus 8): N-rank == 1-rank parity at N = 2 and 8, the bench at N = 8 (phase
# breakdown by rank), BASELINE configs[3] (13B, 3-bit act-order, 40 blocks) and the 7B model at N = 8
tag=${1:-r02_multi8}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
tr() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 "$@"; }
for N in 2 8; do
  timeout 600 bash -c "$(declare -f tr); tr $N tools/multi_gpu_parity.py" > gpurun_out/${tag}_parity_n${N}.json 2> gpurun_out/${tag}_parity_n${N}.err; echo "parity N=$N rc=$?"
  tail -c 600 gpurun_out/${tag}_parity_n${N}.json
done
timeout 600 bash -c "$(declare -f tr); tr 8 bench.py --gpus 8 --steps 5 --warmup 3" > gpurun_out/${tag}_bench_n8.json 2> gpurun_out/${tag}_bench_n8.err; echo "bench N=8 rc=$?"
timeout 900 bash -c "$(declare -f tr); tr 8 tools/full_model.py --model 13b --bits 3" > gpurun_out/${tag}_full_model_13b_3bit_n8.json 2> gpurun_out/${tag}_full_model_13b_3bit_n8.err; echo "13b rc=$?"; tail -c 700 gpurun_out/${tag}_full_model_13b_3bit_n8.json
timeout 900 bash -c "$(declare -f tr); tr 8 tools/full_model.py --model 7b" > gpurun_out/${tag}_full_model_7b_n8.json 2> gpurun_out/${tag}_full_model_7b_n8.err; echo "7b rc=$?"; tail -c 700 gpurun_out/${tag}_full_model_7b_n8.json
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${tag}_bench_n8.json').read().strip().splitlines()[-1])
    print('N', d['n_gpus'], 'value', round(d['value'],5), 'e2e', round(d['e2e']['value'],4), {k:(round(v,5) if isinstance(v,float) else v) for k,v in d['detail'].items() if k.endswith('per_layer')})
    print(d['detail'].get('phase_s_by_rank'))
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/${tag}_bench_n8.err').read()[-2000:])
PY
