#!/usr/bin/env python
"""gpurun_out/launches_<wl>.csv + gpurun_out/prof_bounce_<wl>.ncu-rep (profiles/run_ncu_scene.sh) ->
profiles/<tag>_<wl>_launches.txt and profiles/<tag>_<wl>_bounce_ncu.txt.
Usage: python profiles/summarize_scene.py r01 c3"""
import collections, csv, io, json, os, subprocess, sys

tag, wl = sys.argv[1], sys.argv[2]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")

rows = [l for l in open(os.path.join(G, f"launches_{wl}.csv")) if not l.startswith("==")]
r = list(csv.DictReader(io.StringIO("".join(rows))))
agg = collections.defaultdict(lambda: [0, 0.0])
for x in r:
    n = x["Kernel Name"].split("(")[0][:70]
    agg[n][0] += 1
    agg[n][1] += float(x["Metric Value"].replace(",", ""))
tot = sum(v[1] for v in agg.values())
with open(os.path.join(P, f"{tag}_{wl}_launches.txt"), "w") as f:
    f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none  (profiles/run_ncu_scene.sh {wl})\n")
    f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
    f.write(f"{'launches':>8} {'total_us':>10} {'avg_us':>9} {'share':>6}  kernel\n")
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{c:8d} {t/1e3:10.1f} {t/1e3/c:9.2f} {100*t/tot:5.1f}%  {n}\n")

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum"]
out = subprocess.run(["ncu", "-i", os.path.join(G, f"prof_bounce_{wl}.ncu-rep"), "--page", "raw", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
with open(os.path.join(P, f"{tag}_{wl}_bounce_ncu.txt"), "w") as f:
    f.write(f"# ncu --set full --clock-control none --import-source on -k regex:bounce_kernel -s 21 -c 7  (profiles/run_ncu_scene.sh {wl})\n")
    for row in rows[2:]:
        f.write(f"--- {row[hdr.index('Kernel Name')][:90]}\n")
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
# DRAM bytes of the captured wave (its bounce launches: depth 0..5 + tail), per ONE sample of every pixel
spw = int(sys.argv[3]) if len(sys.argv) > 3 else {"c3": 8, "c4": 4}[wl]      # samples per pixel of the profiled wave (run_ncu_scene.sh SPP)
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tot_b = 0.0
for row in rows[2:]:
    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        tot_b += float(row[hdr.index(k)].replace(",", "")) * scale[units[hdr.index(k)]]
tpath = os.path.join(P, "traffic.json")
tr = json.load(open(tpath)) if os.path.exists(tpath) else {}
tr[f"bounce_kernel:{wl}"] = round(tot_b / spw)
tr["_note"] = ("dram__bytes_read.sum + dram__bytes_write.sum from ncu --set full. csmain_accumulate: per launch (profiles/r01_accumulate_ncu.txt). "
               "bounce_kernel:<workload>: summed over the bounce launches of ONE wave (depth 0..5 + tail) and divided by the wave's samples "
               "per pixel (c3: 8, c4: 4), i.e. per one sample of every pixel = the unit of roofline.achieved "
               "(profiles/r01_c3_bounce_ncu.txt, r01_c4_bounce_ncu.txt)")
json.dump(tr, open(tpath, "w"), indent=1)
print(open(os.path.join(P, f"{tag}_{wl}_launches.txt")).read())
print(open(os.path.join(P, f"{tag}_{wl}_bounce_ncu.txt")).read())
