#!/bin/bash
# K6 windowed mid-batch kernel: parity tests, then the batch sweep through it (W4_PATH=4: wherever it supports the shape),
# through the automatic dispatch (0) and through the dispatch before it existed (3)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pack_w4a16.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/midm_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/midm_tests.log
tail -5 gpurun_out/midm_tests.log
for path in 4 0 3; do
W4_MS=${W4_MS:-3,4,8,9,12,16,17,32} W4_PATH=$path timeout 600 python tools/kernel_bench.py w4a16 > gpurun_out/midm_sweep_path$path.jsonl 2> gpurun_out/midm_sweep_path$path.err
done
python - <<'PY'
import json
def rd(p): return { (d['M'],d['N'],d['K']):d for d in map(json.loads,open('gpurun_out/midm_sweep_path%d.jsonl'%p)) if d.get('k')=='w4a16'}
a,b,c=rd(4),rd(0),rd(3)
for k in a:
    print(k, 'streamed %.1f us  auto %.1f us  before %.1f us  cublas %.1f us  frac_hbm(streamed) %.3f' % (a[k]['us'], b[k]['us'], c[k]['us'], a[k]['dense_cublas_us'], a[k]['frac_hbm']))
PY
