#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sdf_gpu.py -q -x > gpurun_out/pytest_sdf3.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_sdf3.log
timeout 900 python bench.py --workload sdf --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf3.json 2> gpurun_out/bench_sdf3.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_sdf3.json; grep -o '"reference_kernel".*' gpurun_out/bench_sdf3.json | cut -c1-200
timeout 600 python bench.py --workload sdf --points 4096 --subdiv 5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf_small3.json 2> gpurun_out/bench_sdf_small3.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_sdf_small3.json; grep -o '"reference_kernel".*' gpurun_out/bench_sdf_small3.json | cut -c1-200
timeout 600 python bench.py --workload sdf --points 300000 --subdiv 3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf_mid3.json 2> gpurun_out/bench_sdf_mid3.err; echo "bench exit $?"; cut -c1-300 gpurun_out/bench_sdf_mid3.json; grep -o '"reference_kernel".*' gpurun_out/bench_sdf_mid3.json | cut -c1-200
