#!/bin/bash
# f2 (Kin / GPIS optimizer variants): all GPU parity tests, smoke, default bench + pregrasp C2 bench (regression check)
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu9.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/pytest_gpu9.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke9.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke9.log
timeout 600 python bench.py --workload pregrasp_c2 --steps 3 --warmup 3 > gpurun_out/bench_pg9.json 2> gpurun_out/bench_pg9.err; echo "bench exit $?"; cat gpurun_out/bench_pg9.json; tail -5 gpurun_out/bench_pg9.err
