#!/bin/bash
# where does the INT8 std kernel's stage time go: timing with the MMAs removed (exp1) and with the k* arithmetic removed (exp2)
mkdir -p gpurun_out
for v in _exp6; do
  export CDX_LIB=$PWD/compliancedex_b200/libcdx$v.so
  echo "=== variant '$v'"
  timeout 600 python bench.py --workload gpis_c3 --queries 1000000 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/bench_i8$v.json 2> gpurun_out/bench_i8$v.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_i8$v.json
done
