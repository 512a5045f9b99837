#!/bin/bash
tag=${1:-r02_v25}
mkdir -p gpurun_out
export PYTHONDONTWRITEBYTECODE=1
P="python tools/prof_targets.py"
full() { name=$1; args=$2; pat=$3; skip=$4; cnt=$5
  $P $args > gpurun_out/${tag}_${name}_plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$pat" -s $skip -c $cnt -f \
      -o gpurun_out/${tag}_${name} $P $args > gpurun_out/${tag}_${name}_ncu.log 2>&1
  echo "$name exit $?"; }
full tf32_far "tf32 9984 4096 1024" "gemm_tf32x3" 1 1
full tf32_deep "tf32 2816 8192 8192" "gemm_tf32x3" 1 1
full tf32_near "tf32 896 3968 128" "gemm_tf32x3" 1 1
full w4a16mma "w4a16mma" "w4a16_mma_kernel" 1 1
