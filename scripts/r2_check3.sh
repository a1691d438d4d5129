#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpis_gpu.py tests/test_pregrasp_gpu.py tests/test_variants_gpu.py tests/test_configs_gpu.py tests/test_scale_parity_gpu.py -x -q -m gpu > gpurun_out/r2c3_pytest.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/r2c3_pytest.log
timeout 600 python bench.py --workload pregrasp_c5 --no-cpu-baseline > gpurun_out/r2c3_c5.json 2> gpurun_out/r2c3_c5.err; echo "c5 exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r2c3_c5.json')); r=d['roofline']; print(d['value'], d['config']['ms_per_iteration'], r['frac'], r['kernel_ms_per_launch'])"
timeout 600 python bench.py --workload pregrasp_c5 --c5-candidates 8192 --no-cpu-baseline > gpurun_out/r2c3_c5_8192.json 2> gpurun_out/r2c3_c5_8192.err; echo "c5/8192 exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r2c3_c5_8192.json')); r=d['roofline']; print(d['value'], d['config']['ms_per_iteration'], r['frac'], r['kernel_ms_per_launch'])"
timeout 600 python bench.py --workload pregrasp_c2 --no-cpu-baseline > gpurun_out/r2c3_c2.json 2> gpurun_out/r2c3_c2.err; echo "c2 exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r2c3_c2.json')); r=d['roofline']; print(d['value'], d['config']['ms_per_iteration'], r['frac'], r['kernel_ms_per_launch'])"
