#!/bin/bash
tag=${1:-r02_v13}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
( timeout 900 python -m pytest tests/test_gpu_pack_w4a16.py tests/test_native_module.py tests/test_gpu_vs_reference_ext.py tests/test_checkpoint_eval.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
grep -E "^(FAILED|ERROR)|passed|failed|^E  " gpurun_out/${tag}_pytest.log | head -20
timeout 300 python tools/kernel_bench.py gemmnt > gpurun_out/${tag}_gemmnt.jsonl 2>&1
python - <<PY
import json
for l in open('gpurun_out/${tag}_gemmnt.jsonl'):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(d['M'],d['N'],d['K'],'us',round(d['us'],1),'TF',round(d['tflops'],0),'vs cublas',round(d['vs_cublas'],3))
PY
