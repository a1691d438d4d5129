#!/bin/bash
# re-entry baseline: all GPU parity tests, smoke, both bench workloads, fit timing
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu5.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu5.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke5.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/smoke5.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_pg5.json 2> gpurun_out/bench_pg5.err; echo "bench exit $?"; cat gpurun_out/bench_pg5.json; tail -5 gpurun_out/bench_pg5.err
timeout 600 python bench.py --workload gpis_c3 --queries 2000000 --steps 2 --warmup 3 > gpurun_out/bench_c3_5.json 2> gpurun_out/bench_c3_5.err; echo "bench exit $?"; cat gpurun_out/bench_c3_5.json; tail -5 gpurun_out/bench_c3_5.err
timeout 900 python scripts/measure_fit.py > gpurun_out/fit5.json 2> gpurun_out/fit5.err; echo "fit exit $?"; cat gpurun_out/fit5.json; tail -3 gpurun_out/fit5.err
