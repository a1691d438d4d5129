#!/bin/bash
# full GPU check of the working tree: parity suite, smoke, default bench (all blocks), reference arm
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/r2_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/r2_smoke.log
timeout 1200 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench exit $?"; cut -c1-1500 gpurun_out/r2_bench.json; tail -5 gpurun_out/r2_bench.err
if [ "$1" = "ref" ]; then
timeout 1200 python bench.py --impl reference > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref exit $?"; cut -c1-800 gpurun_out/r2_bench_ref.json; tail -3 gpurun_out/r2_bench_ref.err
fi
