"""Instruction census of the built objects (cuobjdump -sass): the SASS mnemonics that prove the Blackwell-native paths -
UTC*MMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), UBLKCP (cp.async.bulk), STAS (st.async),
DMMA (mma.sync f64), LDGSTS (cp.async), SYNCS (mbarrier) - per object file and per kernel.
    python scripts/sass_census.py > profiles/r02_sass_census.json"""
import collections
import glob
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WATCH = ["UTCIMMA", "UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTCCP", "UBLKCP", "UTMALDG", "STAS", "DMMA", "HMMA", "LDGSTS", "SYNCS",
         "MUFU.RSQ", "DFMA", "DMUL", "DADD", "MEMBAR.ALL.GPU", "UCGABAR", "MAPA", "ELECT"]
out = {}
for obj in sorted(glob.glob(os.path.join(ROOT, "compliancedex_b200", "csrc", "build", "*.o"))):
    sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    per_kernel, name = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()[:140]
            per_kernel[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and name:
            op = m.group(1)
            for w in WATCH:
                if op.startswith(w):
                    per_kernel[name][w] += 1
    tot = collections.Counter()
    for c in per_kernel.values():
        tot.update(c)
    out[os.path.basename(obj)] = {"total": dict(tot), "kernels": {k: dict(v) for k, v in per_kernel.items() if v}}
print(json.dumps(out, indent=1))
