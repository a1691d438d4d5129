#!/bin/bash
# sample SM clock / power while the INT8 std kernel runs (2M queries, N=4096)
mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active --format=csv,noheader -lms 100 > gpurun_out/i8_clk_$1.txt &
SMI=$!
CDX_I8_DBG=$1 timeout 300 python scripts/i8_check.py 4096 2000000 > gpurun_out/i8_clkrun_$1.json 2>&1
kill $SMI
sort gpurun_out/i8_clk_$1.txt | uniq -c | sort -rn | head -8
python -c "
import json;d=json.load(open('gpurun_out/i8_clkrun_$1.json'));print($1,[(k,round(v['ms'],1)) for k,v in d.items() if isinstance(v,dict)])"
