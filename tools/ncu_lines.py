"""Per-source-line warp-stall samples from an ncu report captured with --import-source on.
usage: python tools/ncu_lines.py report.ncu-rep [top_n] [extra ncu filters, e.g. -k regex:update_seg -c 1]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
extra = sys.argv[3:]   # e.g. -k regex:update_seg -c 1
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"] + extra,
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; agg = {}
for r in rows:
    if r and r[0] == "Line No":
        hdr = r; ins = hdr.index("# Samples")
        stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        continue
    if hdr is None or len(r) <= ins or not r[0].isdigit():
        continue
    try:
        n = int(r[ins] or 0)
    except ValueError:
        continue
    key = (int(r[0]), r[1].strip()[:90])
    a = agg.setdefault(key, [0, {}])
    a[0] += n
    for j in stalls:
        v = r[j]
        if v.isdigit() and int(v) > 0:
            a[1][hdr[j][6:]] = a[1].get(hdr[j][6:], 0) + int(v)
tot = sum(a[0] for a in agg.values())
print("total samples", tot)
for (ln, src), (n, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    top = sorted(st.items(), key=lambda x: -x[1])[:3]
    print(f"{ln:5d} {n:6d} {100.0*n/tot:5.1f}%  {src}  {top}")
