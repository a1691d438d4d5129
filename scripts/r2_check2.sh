#!/bin/bash
# round 2: first run of the rows-as-M INT8 kernel (bounded by timeouts: a hang must not hold the box)
mkdir -p gpurun_out
timeout 120 python scripts/i8x_check.py 2 > gpurun_out/r2c2_i8x_v2.log 2>&1; echo "variant 2 (no cluster) exit $?"; tail -12 gpurun_out/r2c2_i8x_v2.log
timeout 120 python scripts/i8x_check.py 1 > gpurun_out/r2c2_i8x_v1.log 2>&1; echo "variant 1 (cluster) exit $?"; tail -12 gpurun_out/r2c2_i8x_v1.log
timeout 120 python scripts/i8x_check.py 0 > gpurun_out/r2c2_i8x_v0.log 2>&1; echo "variant 0 (round 1) exit $?"; tail -4 gpurun_out/r2c2_i8x_v0.log
timeout 100 compliancedex_b200/csrc/build/cdx_pipebench > gpurun_out/r2c2_pipebench.json 2>&1; echo "pipebench exit $?"; cat gpurun_out/r2c2_pipebench.json
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
