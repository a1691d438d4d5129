#!/bin/bash
# ncu evidence at the C5 per-GPU shape (Allegro, hammer N=317, 8,192 candidates): launch list + full capture of the iteration's kernels
set -x
mkdir -p gpurun_out
CMD="python bench.py --hand allegro --object hammer --candidates 8192 --steps 1 --warmup 3 --iters 10 --no-cpu-baseline"
$CMD > gpurun_out/plain_c5.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_pregrasp_c5.csv $CMD > gpurun_out/ncu_c5_launches.log 2>&1
echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:"gpis_var_kernel|pg_step_kernel|gpis_mean_kernel" -s 60 -c 6 -f -o gpurun_out/r01_prof_pregrasp_c5 $CMD > gpurun_out/ncu_c5_full.log 2>&1
echo "full exit $?"
