#!/bin/bash
# ncu evidence for the default bench (pregrasp C2) at the current kernels (16-lane step kernel, two candidate groups,
# padded-column MMA skip): launch list + full capture of the three kernels of one iteration.  Runs after check17.
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --iters 10 --no-cpu-baseline"
$CMD > gpurun_out/plain_pg4.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01_launches_pregrasp_v6.csv $CMD > gpurun_out/ncu_pg_launches4.log 2>&1
echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:"gpis_var_kernel|pg_step_kernel|gpis_mean_kernel" -s 60 -c 6 -f -o gpurun_out/r01_prof_pregrasp_v6 $CMD > gpurun_out/ncu_pg_full4.log 2>&1
echo "full exit $?"
