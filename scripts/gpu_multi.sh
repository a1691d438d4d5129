#!/bin/bash
# multi-GPU: bench under torchrun at N=2 (pregrasp default workload + gpis_c3), both arms
set -x
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_pg_n$N.json 2> gpurun_out/bench_pg_n$N.err; echo "exit $?"; cat gpurun_out/bench_pg_n$N.json; tail -5 gpurun_out/bench_pg_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 1 --impl reference > gpurun_out/bench_pg_ref_n$N.json 2> gpurun_out/bench_pg_ref_n$N.err; echo "exit $?"; cat gpurun_out/bench_pg_ref_n$N.json; tail -5 gpurun_out/bench_pg_ref_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload gpis_c3 --queries 2000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "exit $?"; cat gpurun_out/bench_c3_n$N.json; tail -5 gpurun_out/bench_c3_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --workload sdf --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sdf_n$N.json 2> gpurun_out/bench_sdf_n$N.err; echo "exit $?"; cut -c1-400 gpurun_out/bench_sdf_n$N.json; tail -3 gpurun_out/bench_sdf_n$N.err
echo "first bytes of the N=$N stdout files:"; head -c 60 gpurun_out/bench_pg_n$N.json; echo; head -c 60 gpurun_out/bench_c3_n$N.json; echo
