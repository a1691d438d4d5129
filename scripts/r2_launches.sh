#!/bin/bash
# ncu launch list (gpu__time_duration per launch; cold-cache and serialised: compare SHARES) of one bench workload
# usage: bash scripts/r2_launches.sh <tag> <bench args...>
tag=$1; shift
mkdir -p gpurun_out
python bench.py "$@" --no-cpu-baseline > gpurun_out/launches_${tag}_plain.json 2> gpurun_out/launches_${tag}_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_COUNT:-600} --csv --log-file gpurun_out/r02_launches_$tag.csv python bench.py "$@" --no-cpu-baseline > gpurun_out/launches_${tag}_ncu.log 2>&1
echo "exit $?"; python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/r02_launches_$tag.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name, val, unit = r[4][:70], float(r[-1].replace(",", "")), r[-2]
    ms = val * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
    agg[name][0] += 1; agg[name][1] += ms
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print("%-72s n=%4d %10.3f ms %5.1f %%" % (k, v[0], v[1], 100 * v[1] / tot))
PY
