#!/bin/bash
# 1 GPU: refresh ncu evidence for the current kernels (default bench launch list; full capture of the SDF pair kernel at the bench size)
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --iters 10 --no-cpu-baseline"
$CMD > gpurun_out/plain_pg12.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_pregrasp_v5.csv $CMD > gpurun_out/ncu_pg12_launches.log 2>&1
echo "launch list exit $?"
CMD2="python bench.py --workload sdf --steps 1 --warmup 3 --no-cpu-baseline"
$CMD2 > gpurun_out/plain_sdf12.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"sdf_pair_kernel" -s 2 -c 1 -f -o gpurun_out/r01_prof_sdf_v2 $CMD2 > gpurun_out/ncu_sdf12.log 2>&1
echo "ncu exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r01_launches_sdf_v2.csv $CMD2 > gpurun_out/ncu_sdf12_launches.log 2>&1
echo "launch list exit $?"
