#!/bin/bash
# pregrasp: parity tests + first bench line (both arms) + launch list
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu3.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu3.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_pg.json 2> gpurun_out/bench_pg.err; echo "bench exit $?"; cat gpurun_out/bench_pg.json; tail -5 gpurun_out/bench_pg.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_pg_ref.json 2> gpurun_out/bench_pg_ref.err; echo "ref exit $?"; cat gpurun_out/bench_pg_ref.json; tail -5 gpurun_out/bench_pg_ref.err
