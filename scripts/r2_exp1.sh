#!/bin/bash
mkdir -p gpurun_out
for v in 1 2; do for e in 0 1 2 3 4 5; do timeout 60 python scripts/i8x_exp.py $e $v 2>&1 | tail -1; done; done | tee gpurun_out/r2_i8x_exp1.log
