#!/bin/bash
# N GPUs of one box at the current kernels: default pregrasp bench (C2, weak scaling), the FULL C5 shape (Allegro, 8,192
# candidates per GPU x 500 iterations = 65,536 candidates on 8 GPUs, NCCL arg-min) for hammer / lego / coffeebottle,
# and gpis_c3 at its full 10 M queries
set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/m8_pg_n$N.json 2> gpurun_out/m8_pg_n$N.err; echo "exit $?"; cut -c1-500 gpurun_out/m8_pg_n$N.json; tail -3 gpurun_out/m8_pg_n$N.err
for o in hammer lego coffeebottle; do
timeout 600 $TR --master-port 29522 bench.py --gpus $N --hand allegro --object $o --candidates 8192 --iters 500 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/m8_c5_${o}_n$N.json 2> gpurun_out/m8_c5_${o}_n$N.err; echo "exit $?"; cut -c1-700 gpurun_out/m8_c5_${o}_n$N.json; tail -3 gpurun_out/m8_c5_${o}_n$N.err
done
timeout 600 $TR --master-port 29523 bench.py --gpus $N --workload gpis_c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/m8_c3_n$N.json 2> gpurun_out/m8_c3_n$N.err; echo "exit $?"; cut -c1-500 gpurun_out/m8_c3_n$N.json; tail -3 gpurun_out/m8_c3_n$N.err
