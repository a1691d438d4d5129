#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sdf_gpu.py tests/test_abi.py -q -x > gpurun_out/pytest_sdf4.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/pytest_sdf4.log
timeout 900 python bench.py --workload sdf --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf4.json 2> gpurun_out/bench_sdf4.err; echo "bench exit $?"; cut -c1-260 gpurun_out/bench_sdf4.json; grep -o '"reference_kernel".*' gpurun_out/bench_sdf4.json | cut -c1-900; tail -3 gpurun_out/bench_sdf4.err
timeout 600 python bench.py --workload sdf --points 4096 --subdiv 5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf_small4.json 2> gpurun_out/bench_sdf_small4.err; echo "bench exit $?"; cut -c1-260 gpurun_out/bench_sdf_small4.json; grep -o '"reference_kernel".*' gpurun_out/bench_sdf_small4.json | cut -c1-900
