#!/bin/bash
# check of the committed code: full GPU parity suite, smoke, default bench, reference arm, sdf + gpis_c3 bench lines
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/final_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/final_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/final_smoke.log
timeout 600 python bench.py > gpurun_out/final_bench_pg.json 2> gpurun_out/final_bench_pg.err; echo "bench exit $?"; cat gpurun_out/final_bench_pg.json; tail -3 gpurun_out/final_bench_pg.err
if [ "$1" = "all" ]; then
timeout 600 python bench.py --impl reference > gpurun_out/final_bench_pg_ref.json 2> gpurun_out/final_bench_pg_ref.err; echo "ref exit $?"; cat gpurun_out/final_bench_pg_ref.json
timeout 600 python bench.py --workload gpis_c3 --no-cpu-baseline > gpurun_out/final_bench_c3.json 2> gpurun_out/final_bench_c3.err; echo "c3 exit $?"; cut -c1-600 gpurun_out/final_bench_c3.json
timeout 600 python bench.py --workload sdf --no-cpu-baseline > gpurun_out/final_bench_sdf.json 2> gpurun_out/final_bench_sdf.err; echo "sdf exit $?"; cut -c1-600 gpurun_out/final_bench_sdf.json
fi
