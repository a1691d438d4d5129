#!/bin/bash
# step kernel with full-warp shuffles + occupancy variants (CDX_PG_OCC): parity suite, C2 bench, C5 per-GPU shape with each variant
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/c19_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/c19_pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/c19_pg.json 2> gpurun_out/c19_pg.err; echo "bench exit $?"
for occ in 2 3 auto; do
for o in hammer coffeebottle; do
if [ $occ = auto ]; then unset CDX_PG_OCC; else export CDX_PG_OCC=$occ; fi
timeout 300 python bench.py --hand allegro --object $o --candidates 8192 --iters 100 --no-cpu-baseline > gpurun_out/c19_c5_${o}_occ$occ.json 2> gpurun_out/c19_c5_${o}_occ$occ.err; echo "bench exit $?"
done
done
python - <<'P'
import json,glob
for f in sorted(glob.glob("gpurun_out/c19_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, "%.4g"%d["value"], d["config"]["ms_per_iteration"], d["roofline"]["kernel_ms_per_iteration"])
    except Exception as e: print(f,"ERR",e)
P
