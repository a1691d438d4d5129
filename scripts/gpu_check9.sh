#!/bin/bash
# f3: SDF tests after pinning the dist FMA pattern, bench --workload sdf (with the reference kernel timed beside), ncu of the pair kernel
set -x
mkdir -p gpurun_out
timeout 600 python scripts/make_golden_sdf.py > gpurun_out/sdf_golden.log 2>&1; echo "golden exit $?"; tail -6 gpurun_out/sdf_golden.log
timeout 900 python -m pytest tests/test_sdf_gpu.py tests/test_sdf_oracle.py -q -x > gpurun_out/pytest_sdf.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_sdf.log
timeout 900 python bench.py --workload sdf --steps 3 --warmup 3 > gpurun_out/bench_sdf1.json 2> gpurun_out/bench_sdf1.err; echo "bench exit $?"; cat gpurun_out/bench_sdf1.json; tail -5 gpurun_out/bench_sdf1.err
timeout 600 python bench.py --workload sdf --points 4096 --subdiv 5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf_small.json 2> gpurun_out/bench_sdf_small.err; echo "bench exit $?"; cat gpurun_out/bench_sdf_small.json
CMD="python bench.py --workload sdf --points 300000 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_sdf.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"sdf_pair_kernel" -s 2 -c 1 -f -o gpurun_out/r01_prof_sdf $CMD > gpurun_out/ncu_sdf.log 2>&1
echo "ncu exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r01_launches_sdf.csv $CMD > gpurun_out/ncu_sdf_launches.log 2>&1
echo "launch list exit $?"
