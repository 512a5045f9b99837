#!/bin/bash
# ncu evidence for the shipped kernels: ordered launch list of one factorisation + one linear, --set full
# captures of the kernels that had no capture of their final form.  Each ncu command runs after the same command exited 0.
tag=${1:-r02_v16}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
P="python tools/prof_targets.py"
full() {  # name, target args, kernel regex, skip, count
  name=$1; args=$2; pat=$3; skip=$4; cnt=$5
  $P $args > gpurun_out/${tag}_${name}_plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$pat" -s $skip -c $cnt -f \
      -o gpurun_out/${tag}_${name} $P $args > gpurun_out/${tag}_${name}_ncu.log 2>&1
  echo "$name exit $?"
}
# ordered launch lists (cold-cache, serialised): where does a factorisation / a linear spend its launches
for a in "invert 4096" "invert 11008" "layer 11008 4096" "layer 4096 11008"; do
  n=$(echo $a | tr ' ' '_')
  $P $a > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv \
      --log-file gpurun_out/${tag}_launches_${n}.csv $P $a > gpurun_out/${tag}_launches_${n}.log 2>&1
  echo "launches $n exit $?"
done
full solve_nq4 "layer 11008 4096" "solve_super_kernel" 1 1
full solve_nq1 "layer 4096 11008" "solve_super_kernel" 1 1
full assembly "layer 11008 4096" "transpose_pack_kernel|prep_weight_scatter_kernel|gather_kernel|reverse_lower_kernel|trtri_diag_kernel|solve_metrics_kernel" 6 6
full chol_panel "invert 4096" "chol_panel_kernel" 40 2
full w4a16mma "w4a16mma" "w4a16" 2 1
full gemmnt "gemmnt" "gemm_f16_nt_kernel" 2 1
ls -la gpurun_out | tail -30
