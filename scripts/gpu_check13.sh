#!/bin/bash
# pregrasp: mean query launched first; rolled finger loops variant
mkdir -p gpurun_out
for v in "" _roll; do
  export CDX_LIB=$PWD/compliancedex_b200/libcdx$v.so
  echo "=== variant '$v'"
  timeout 900 python -m pytest tests/test_pregrasp_gpu.py tests/test_variants_gpu.py tests/test_configs_gpu.py -q -x 2>&1 | tail -2
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pg13$v.json 2> gpurun_out/bench_pg13$v.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_pg13$v.json')); print(d['value'], d['config']['ms_per_iteration'], d['e2e']['value'], d['roofline']['kernel_ms_per_iteration'])"
  timeout 600 python bench.py --hand allegro --object hammer --candidates 8192 --iters 100 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/bench_c5s13$v.json 2> gpurun_out/bench_c5s13$v.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_c5s13$v.json')); print(d['value'], d['config']['ms_per_iteration'], d['roofline']['kernel_ms_per_iteration'])"
done
