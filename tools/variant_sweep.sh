#!/bin/bash
# A/B of library variants built into gptq-v2_b200/lib/var_<name>.so: swaps each in (on the GPU box's copy only) and runs a
# K6 sweep.  usage: variant_sweep.sh "<names>" ; env W4_MS, W4_SHAPES as tools/kernel_bench.py
mkdir -p gpurun_out
L=gptq-v2_b200/lib
cp $L/libgptq_b200.so /tmp/lib_orig.so
for v in $1; do
  cp $L/var_$v.so $L/libgptq_b200.so
  timeout 600 python tools/kernel_bench.py w4a16 > gpurun_out/var_$v.jsonl 2> gpurun_out/var_$v.err || tail -3 gpurun_out/var_$v.err
done
cp /tmp/lib_orig.so $L/libgptq_b200.so
python - "$1" <<'PY'
import json, sys
names = sys.argv[1].split()
r = {n: {(d['M'], d['N'], d['K']): d for d in map(json.loads, open('gpurun_out/var_%s.jsonl' % n)) if d.get('k') == 'w4a16'} for n in names}
for k in r[names[0]]:
    print(k, '  '.join('%s %.1f' % (n, r[n][k]['us']) for n in names if k in r[n]))
PY
