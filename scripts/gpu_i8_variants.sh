#!/bin/bash
# INT8 std kernel scheduling experiments: BK=32 / 4 TMEM stages / generator warps split over two stages, and completing a
# stage's tcgen05.st after the next stage has been generated
mkdir -p gpurun_out
for v in "" _bk32 _late _late32; do
  export CDX_LIB=$PWD/compliancedex_b200/libcdx$v.so
  echo "=== variant '$v'"
  timeout 600 python -m pytest tests/test_gpis_i8_gpu.py -q -x 2>&1 | tail -2
  timeout 600 python bench.py --workload gpis_c3 --queries 2000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_i8var$v.json 2> gpurun_out/bench_i8var$v.err; echo "bench exit $?"; cut -c1-200 gpurun_out/bench_i8var$v.json
done
