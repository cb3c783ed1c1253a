#!/usr/bin/env python
"""Turns gpurun_out/launches.csv + gpurun_out/prof_accumulate.ncu-rep (profiles/run_ncu.sh) into the
tracked summaries profiles/rNN_launches.txt, profiles/rNN_accumulate_ncu.txt and profiles/traffic.json.
Usage: python profiles/summarize.py r01"""
import collections, csv, io, json, os, subprocess, sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")

rows = [l for l in open(os.path.join(G, "launches.csv")) if not l.startswith("==")]
r = list(csv.DictReader(io.StringIO("".join(rows))))
agg = collections.defaultdict(lambda: [0, 0.0])
for x in r:
    n = x["Kernel Name"].split("(")[0][:70]
    agg[n][0] += 1
    agg[n][1] += float(x["Metric Value"].replace(",", ""))
tot = sum(v[1] for v in agg.values())
with open(os.path.join(P, f"{tag}_launches.txt"), "w") as f:
    f.write("# ncu --metrics gpu__time_duration.sum --clock-control none  python bench.py --steps 5 --warmup 3 --no-cpu-baseline\n")
    f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
    f.write(f"{'launches':>8} {'total_us':>10} {'avg_us':>9} {'share':>6}  kernel\n")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{c:8d} {t/1e3:10.1f} {t/1e3/c:9.2f} {100*t/tot:5.1f}%  {n}\n")

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg"]
out = subprocess.run(["ncu", "-i", os.path.join(G, "prof_accumulate.ncu-rep"), "--page", "raw", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
traffic = {}
with open(os.path.join(P, f"{tag}_accumulate_ncu.txt"), "w") as f:
    f.write("# ncu --set full --clock-control none --import-source on -k regex:csmain_accumulate -s 3 -c 2  (same bench command)\n")
    for row in rows[2:]:
        name = row[hdr.index("Kernel Name")]
        f.write(f"--- {name}\n")
        for k in KEYS:
            if k in hdr:
                f.write(f"  {k} = {row[hdr.index(k)]} {units[hdr.index(k)]}\n")
        stalls = []
        for i, k in enumerate(hdr):
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
                try:
                    stalls.append((float(row[i]), k))
                except ValueError:
                    pass
        for v, k in sorted(stalls, reverse=True)[:6]:
            f.write(f"  {k} = {v:.3f}\n")
        def val(k):
            v, u = float(row[hdr.index(k)].replace(",", "")), units[hdr.index(k)]
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        traffic["csmain_accumulate"] = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
tpath = os.path.join(P, "traffic.json")
merged = json.load(open(tpath)) if os.path.exists(tpath) else {}      # keep the entries of the scene workloads
merged.update(traffic)
json.dump(merged, open(tpath, "w"), indent=1)
print(open(os.path.join(P, f"{tag}_launches.txt")).read())
print(open(os.path.join(P, f"{tag}_accumulate_ncu.txt")).read())
