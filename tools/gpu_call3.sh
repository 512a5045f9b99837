#!/bin/bash
tag=${1:-r02_v3}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
( timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
tail -3 gpurun_out/${tag}_pytest.log
for s in 1 2 3 4; do GQ_SUPERS=$s timeout 120 python tools/kernel_bench.py solve; done > gpurun_out/${tag}_solve_fused.jsonl 2>&1
GQ_SOLVE_FUSED=0 GQ_SUPERS=8 timeout 120 python tools/kernel_bench.py solve > gpurun_out/${tag}_solve_unfused.jsonl 2>&1
cat gpurun_out/${tag}_solve_fused.jsonl gpurun_out/${tag}_solve_unfused.jsonl
timeout 120 python tools/kernel_bench.py invert > gpurun_out/${tag}_invert.jsonl 2>&1; cat gpurun_out/${tag}_invert.jsonl
timeout 600 python tools/parity_probe.py 11008 4 256 > gpurun_out/${tag}_probe_11008.json 2> gpurun_out/${tag}_probe_11008.err; echo "probe rc=$?"; cat gpurun_out/${tag}_probe_11008.json
timeout 300 python tools/parity_probe.py 4096 4 256 > gpurun_out/${tag}_probe_4096.json 2> gpurun_out/${tag}_probe_4096.err; echo "probe rc=$?"; cat gpurun_out/${tag}_probe_4096.json
bash tools/profile_round.sh ${tag} launches
