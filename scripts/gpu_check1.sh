#!/bin/bash
# first GPU pass: pipe peaks, GPIS parity tests, short bench, ncu launch list
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
compliancedex_b200/csrc/build/cdx_pipebench > gpurun_out/pipebench.json 2>&1; cat gpurun_out/pipebench.json
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload gpis_c3 --queries 2000000 --steps 2 --warmup 3 > gpurun_out/bench_gpis.json 2> gpurun_out/bench_gpis.err; echo "bench exit $?"; cat gpurun_out/bench_gpis.json; tail -5 gpurun_out/bench_gpis.err
