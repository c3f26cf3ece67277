"""Per-kernel counts of the Blackwell-specific SASS instructions in libb200km.so (tcgen05 MMA / TMEM / TMA / mbarrier),
written as a markdown table.  usage: python tools/sass_digest.py [path/to/libb200km.so] > profiles/rN_sass_digest.md
Command underneath: cuobjdump -sass <lib>"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "dask_ml_b200", "libb200km.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
WATCH = ["UTCHMMA", "UTCBAR", "LDTM", "UTCATOMSWS", "UBLKCP", "UTMALDG", "SYNCS", "FMNMX3", "FFMA.SAT", "DFMA", "DADD", "FFMA",
         "HMMA", "DMMA", "LDGSTS", "ATOMG", "RED", "BAR.SYNC", "UCGABAR"]
kern = None
counts = collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern)
        counts[kern] = collections.Counter()
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        counts[kern]["_total"] += 1
        for w in WATCH:
            if op == w or op.startswith(w + "."):
                counts[kern][w] += 1
print("# SASS digest of `dask_ml_b200/libb200km.so` (`cuobjdump -sass`, sm_100a) — instruction counts per kernel\n")
print("`UTCHMMA` = tcgen05.mma, `LDTM` = tcgen05.ld (TMEM -> registers), `UTCBAR` = tcgen05.commit, `UBLKCP` = "
      "cp.async.bulk (1-D TMA), `SYNCS` = mbarrier ops, `FMNMX3` = three-input min/max, `UCGABAR` = cluster barrier.\n")
cols = [w for w in WATCH if any(c[w] for c in counts.values())]
print("| kernel | instructions | " + " | ".join(cols) + " |")
print("|---|---|" + "---|" * len(cols))
for k, c in counts.items():
    if c["_total"] == 0:
        continue
    print(f"| `{k}` | {c['_total']} | " + " | ".join(str(c[w]) if c[w] else "" for w in cols) + " |")
