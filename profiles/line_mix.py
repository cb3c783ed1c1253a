#!/usr/bin/env python
"""Per-source-line executed warp-instructions and active threads from an .ncu-rep (needs -lineinfo + --import-source on).
Usage: python profiles/line_mix.py report.ncu-rep [top [launch-skip]]   (one launch: the one after `launch-skip` others)"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
sel = ["--launch-skip", sys.argv[3], "--launch-count", "1"] if len(sys.argv) > 3 else []
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"] + sel, capture_output=True, text=True).stdout
lines = out.split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith('"Line No"'))
rows = list(csv.reader(io.StringIO("\n".join(lines[start:]))))
h = rows[0]
iline, isrc = 0, 1
ii, it = h.index("Instructions Executed"), h.index("Thread Instructions Executed")
agg = collections.OrderedDict()
cur = None
fname = ""
for r in rows[1:]:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if len(r) <= it or not (r[iline].isdigit() or r[iline] == ""): continue
    if r[iline]:
        cur = (fname + ":" + r[iline], r[isrc].strip())
    if cur is None or not r[ii].isdigit(): continue
    a = agg.setdefault(cur, [0, 0])
    a[0] += int(r[ii]); a[1] += int(r[it])
tot = sum(v[0] for v in agg.values())
print(f"total {tot} warp-instructions, {sum(v[1] for v in agg.values())/tot:.1f} threads/inst")
for (ln, src), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{ln:>22s} {100*n/tot:5.1f}% thr {t/max(n,1):5.1f}  {src[:110]}")
