#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_configs_gpu.py -m gpu -q > gpurun_out/pytest_gpu4.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/pytest_gpu4.log
timeout 900 python scripts/measure_fit.py > gpurun_out/fit.json 2> gpurun_out/fit.err; echo "fit exit $?"; cat gpurun_out/fit.json; tail -3 gpurun_out/fit.err
