#!/bin/bash
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --iters 10 --no-cpu-baseline"
$CMD > gpurun_out/plain_pg2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gpis_var_kernel|pg_step_kernel" -s 40 -c 4 -o gpurun_out/r01_prof_pregrasp_v3 $CMD > gpurun_out/ncu_pg_full2.log 2>&1
echo "exit $?"
