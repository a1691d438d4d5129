#!/bin/bash
# step kernel: section ticks of the current kernel + ncu with and without cache flushes (is "no_instruction" an artefact
# of ncu's cold instruction cache?); refreshed ncu evidence of the committed kernels
set -x
mkdir -p gpurun_out
CDX_LIB=$PWD/compliancedex_b200/libcdx_ticks.so timeout 300 python scripts/pg_ticks.py 1024 2>&1 | tail -1 | tee gpurun_out/pg_ticks_v6.json
bash scripts/gpu_ncu4.sh
CMD="python bench.py --steps 1 --warmup 3 --iters 10 --no-cpu-baseline"
ncu --set full --clock-control none --cache-control none --import-source on -k regex:"pg_step_kernel" -s 20 -c 2 -f -o gpurun_out/r01_prof_pg_step_warm $CMD > gpurun_out/ncu_pg_warm.log 2>&1
echo "warm exit $?"
