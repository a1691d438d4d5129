#!/bin/bash
mkdir -p gpurun_out
CDX_LIB=$PWD/compliancedex_b200/libcdx_ticks.so timeout 300 python scripts/pg_ticks.py 1024 2>&1 | tail -2 | tee gpurun_out/pg_ticks.json
for v in ""; do
  export CDX_LIB=$PWD/compliancedex_b200/libcdx$v.so
  echo "=== variant '$v'"
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2> gpurun_out/v.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['config']['ms_per_iteration'], d['roofline']['kernel_ms_per_iteration'])"
done
unset CDX_LIB
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
