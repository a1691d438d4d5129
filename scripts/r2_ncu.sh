#!/bin/bash
# one ncu --set full capture per call (profiling recipe): bash scripts/r2_ncu.sh <tag> <kernel regex> <skip> <count> <target args...>
tag=$1; regex=$2; skip=$3; count=$4; shift 4
mkdir -p gpurun_out
python scripts/ncu_target.py "$@" > gpurun_out/ncu_${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c $count -f -o gpurun_out/r02_ncu_$tag python scripts/ncu_target.py "$@" > gpurun_out/ncu_${tag}.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/ncu_${tag}.log; ls -la gpurun_out/r02_ncu_$tag.ncu-rep
