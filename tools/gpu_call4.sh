#!/bin/bash
tag=${1:-r02_v4}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
( timeout 900 python -m pytest tests/test_gpu_solver.py tests/test_gpu_e2e.py tests/test_gpu_quantizer.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${tag}_pytest.log )
tail -3 gpurun_out/${tag}_pytest.log
GQ_SUPERS=4 timeout 120 python tools/kernel_bench.py solve > gpurun_out/${tag}_solve_fused.jsonl 2>&1
GQ_SOLVE_FUSED=0 GQ_SUPERS=8 timeout 120 python tools/kernel_bench.py solve > gpurun_out/${tag}_solve_unfused.jsonl 2>&1
cat gpurun_out/${tag}_solve_fused.jsonl gpurun_out/${tag}_solve_unfused.jsonl
timeout 300 python tools/parity_probe.py 4096 factor > gpurun_out/${tag}_factor_4096.json 2>&1; cat gpurun_out/${tag}_factor_4096.json
timeout 600 python tools/parity_probe.py 11008 factor > gpurun_out/${tag}_factor_11008.json 2>&1; cat gpurun_out/${tag}_factor_11008.json
bash tools/profile_round.sh ${tag} full "solve" solve_super
python tools/ncu_summary.py --stalls gpurun_out/${tag}_solve.ncu-rep > gpurun_out/${tag}_ncu_solve_super_summary.txt 2>&1; cat gpurun_out/${tag}_ncu_solve_super_summary.txt
