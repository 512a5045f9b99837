#!/bin/bash
tag=${1:-r02_v11}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
export GQ_SHAPES=4096x4096,11008x4096,4096x11008,5120x5120,13824x5120,5120x13824
GQ_SUPERS=2,3,4 timeout 300 python tools/kernel_bench.py solve > gpurun_out/${tag}_solve_fused.jsonl 2>&1
GQ_SOLVE_FUSED=0 GQ_SUPERS=4,8 timeout 300 python tools/kernel_bench.py solve > gpurun_out/${tag}_solve_unfused.jsonl 2>&1
cat gpurun_out/${tag}_solve_fused.jsonl gpurun_out/${tag}_solve_unfused.jsonl | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(d['R'],d['C'],'fused',d.get('fused'),'super',d['super'],'ms',round(d['ms'],3),'us/blk',round(d['us_per_block'],1))
"
