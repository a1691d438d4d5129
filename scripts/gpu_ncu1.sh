#!/bin/bash
# ncu evidence, round 1: launch list of a short pregrasp bench + full capture of the two dominant kernels; GPIS C3 unit kernel
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --iters 10 --no-cpu-baseline"
$CMD > gpurun_out/plain_pg.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 300 --csv --log-file gpurun_out/r01_launches_pregrasp.csv $CMD > gpurun_out/ncu_pg_launches.log 2>&1
echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:"gpis_var_kernel|pg_objective_kernel|pg_points_kernel|gpis_mean_kernel" -s 60 -c 4 -o gpurun_out/r01_prof_pregrasp $CMD > gpurun_out/ncu_pg_full.log 2>&1
echo "full pregrasp exit $?"
CMD2="python bench.py --workload gpis_c3 --queries 262144 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD2 > gpurun_out/plain_c3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gpis_var_kernel" -s 3 -c 1 -o gpurun_out/r01_prof_gpis_c3 $CMD2 > gpurun_out/ncu_c3_full.log 2>&1
echo "full c3 exit $?"
ls -la gpurun_out
