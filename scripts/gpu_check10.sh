#!/bin/bash
# branch-free SDF region select + templated pregrasp variants: all GPU tests, sdf + pregrasp benches
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu10.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/pytest_gpu10.log
timeout 900 python bench.py --workload sdf --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf2.json 2> gpurun_out/bench_sdf2.err; echo "bench exit $?"; cat gpurun_out/bench_sdf2.json; tail -5 gpurun_out/bench_sdf2.err
timeout 600 python bench.py --workload sdf --points 4096 --subdiv 5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf_small2.json 2> gpurun_out/bench_sdf_small2.err; echo "bench exit $?"; cat gpurun_out/bench_sdf_small2.json
timeout 600 python bench.py --workload pregrasp_c2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pg10.json 2> gpurun_out/bench_pg10.err; echo "bench exit $?"; cat gpurun_out/bench_pg10.json; tail -5 gpurun_out/bench_pg10.err
