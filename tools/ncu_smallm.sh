set -x
mkdir -p gpurun_out
P="python tools/prof_targets.py"
$P w4a16 1 28672 8192 > gpurun_out/smallm_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k "regex:w4a16_smallm_kernel" -s 2 -c 1 -f -o gpurun_out/r02_v44_smallm_m1 $P w4a16 1 28672 8192 > gpurun_out/smallm_ncu.log 2>&1
echo exit $?
ls -la gpurun_out/*.ncu-rep
