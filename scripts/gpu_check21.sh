#!/bin/bash
# final sanity of the committed code + candidate-group count at the C5 per-GPU shape (CDX_PG_GROUPS 1..4)
set -x
mkdir -p gpurun_out
bash scripts/gpu_final.sh
for G in 1 2 3 4; do
for o in hammer coffeebottle; do
CDX_PG_GROUPS=$G timeout 300 python bench.py --hand allegro --object $o --candidates 8192 --iters 100 --no-cpu-baseline > gpurun_out/c21_c5_${o}_g$G.json 2> gpurun_out/c21_c5_${o}_g$G.err; echo "bench exit $?"
done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/c21_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "%.4g"%d["value"], "%.4f"%d["config"]["ms_per_iteration"], d["config"]["argmin"])
    except Exception as e: print(f,"ERR",e)
P
