#!/bin/bash
# round 2, first GPU check: new host loop (CUDA graph, tuning ABI, NaN record, ownership), new GPIS parity tests, scale parity stats
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_pregrasp_gpu.py tests/test_variants_gpu.py -x -q -m gpu > gpurun_out/r2c1_pytest_pg.log 2>&1; echo "pytest pg exit $?"; tail -15 gpurun_out/r2c1_pytest_pg.log
timeout 900 python -m pytest tests/test_gpis_gpu.py tests/test_gpis_i8_gpu.py -x -q -m gpu > gpurun_out/r2c1_pytest_gpis.log 2>&1; echo "pytest gpis exit $?"; tail -15 gpurun_out/r2c1_pytest_gpis.log
timeout 600 python scripts/scale_parity.py c2 --oracle 1024 > gpurun_out/r2c1_scale_c2.log 2>&1; echo "scale c2 exit $?"; tail -5 gpurun_out/r2c1_scale_c2.log
timeout 600 python scripts/scale_parity.py c5 > gpurun_out/r2c1_scale_c5.log 2>&1; echo "scale c5 exit $?"; tail -5 gpurun_out/r2c1_scale_c5.log
timeout 300 python bench.py --workload pregrasp_c2 --no-cpu-baseline > gpurun_out/r2c1_bench_c2.json 2> gpurun_out/r2c1_bench_c2.err; echo "bench exit $?"; cut -c1-400 gpurun_out/r2c1_bench_c2.json; tail -3 gpurun_out/r2c1_bench_c2.err
nproc
