#!/bin/bash
# INT8 std path: all GPU parity tests, gpis_c3 bench (auto = INT8, and FP64 MMA for comparison), ncu capture
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu6.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu6.log
timeout 600 python bench.py --workload gpis_c3 --queries 2000000 --steps 3 --warmup 3 > gpurun_out/bench_c3_i8.json 2> gpurun_out/bench_c3_i8.err; echo "bench exit $?"; cat gpurun_out/bench_c3_i8.json; tail -5 gpurun_out/bench_c3_i8.err
timeout 600 python bench.py --workload gpis_c3 --queries 2000000 --steps 2 --warmup 3 --var-mode -1 --no-cpu-baseline > gpurun_out/bench_c3_dmma.json 2> gpurun_out/bench_c3_dmma.err; echo "bench exit $?"; cat gpurun_out/bench_c3_dmma.json
CMD="python bench.py --workload gpis_c3 --queries 37888 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_c3_i8.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gpis_std_i8_kernel" -s 2 -c 1 -f -o gpurun_out/r01_prof_gpis_c3_i8 $CMD > gpurun_out/ncu_c3_i8.log 2>&1
echo "ncu exit $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r01_launches_gpis_c3_i8.csv $CMD > gpurun_out/ncu_c3_launches.log 2>&1
echo "launch list exit $?"
