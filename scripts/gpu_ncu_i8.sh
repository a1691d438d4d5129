#!/bin/bash
# ncu full capture of the tcgen05 INT8 std kernel (N=4096, 37,888 queries = 2 waves)
set -x
mkdir -p gpurun_out
CMD="python scripts/i8_check.py 4096 37888"
$CMD > gpurun_out/plain_i8.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"gpis_std_i8_kernel" -s 4 -c 1 -o gpurun_out/r01_prof_i8b -f $CMD > gpurun_out/ncu_i8.log 2>&1
echo "exit $?"; tail -3 gpurun_out/ncu_i8.log
