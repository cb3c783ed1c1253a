#!/usr/bin/env python
"""Turns the round-2 captures in gpurun_out/ (profiles/run_ncu_r02.sh) into the tracked summaries:
   profiles/r02_launches_c2.txt, r02_c2_accumulate_ncu.txt, r02_c2_literal_ncu.txt, r02_c3_bounce_ncu.txt,
   r02_c4_bounce_ncu.txt, r02_peer_merge_ncu.txt and profiles/traffic.json.  Missing captures are skipped.
Usage: python profiles/summarize_r02.py"""
import collections, csv, io, json, os, subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6,
        "ns": 1e-3, "us": 1.0, "ms": 1e3}
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "lts__t_sectors_srcunit_tex_op_read.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "l1tex__t_sector_hit_rate.pct",
        "lts__t_sector_hit_rate.pct", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "nvlrx__bytes.sum", "nvltx__bytes.sum"]


def load(rep):
    path = os.path.join(G, rep)
    if not os.path.exists(path):
        return None
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    res = []
    for row in rows[2:]:
        d = {"_name": row[hdr.index("Kernel Name")], "_device": row[hdr.index("Device")] if "Device" in hdr else ""}
        for i, k in enumerate(hdr):
            try:
                d[k] = (float(row[i].replace(",", "")), units[i])
            except ValueError:
                pass
        res.append(d)
    return res


def val(d, k):
    """metric in base units (bytes, microseconds, plain numbers)"""
    if k not in d:
        return None
    v, u = d[k]
    return v * UNIT.get(u, 1)


def short(name):
    return name.replace("void ptb::", "").replace("<unnamed>::", "").split("(ptb::")[0].split("(const ")[0][:80]


def write_kernels(fn, header, kernels):
    with open(os.path.join(P, fn), "w") as f:
        f.write(header)
        for d in kernels:
            f.write(f"--- {short(d['_name'])}" + (f"   [device {d['_device']}]" if d["_device"] else "") + "\n")
            for k in KEYS:
                if k in d:
                    f.write(f"  {k} = {d[k][0]:,.6g} {d[k][1]}\n")
            stalls = sorted(((d[k][0], k) for k in d if "issue_stalled" in k and k.endswith("per_issue_active.ratio")), reverse=True)
            for v, k in stalls[:6]:
                f.write(f"  {k} = {v:.3f}\n")


traffic = {"_note": "per-kernel DRAM / L2 / instruction figures from ncu --set full (round 2, profiles/run_ncu_r02.sh); bench.py reads "
                    "csmain_accumulate.dram_bytes_per_launch (-> roofline.traffic) and warp_inst_per_sample / "
                    "active_threads_per_inst (-> issue-slot rooflines).  DRAM writes are reported separately: the accumulation "
                    "write-back stays in the 126 MB L2 while the kernel runs (l2_write_bytes_per_launch is what the L2 "
                    "received) and reaches DRAM later, by eviction."}
old = os.path.join(P, "traffic.json")
if os.path.exists(old):
    try:
        prev = json.load(open(old))
        for k, v in prev.items():
            if isinstance(v, dict):
                traffic[k] = v
    except Exception:
        pass

# ---- c2 launch list --------------------------------------------------------------------------------------------------
lp = os.path.join(G, "r02_launches_c2.csv")
if os.path.exists(lp):
    rows = [l for l in open(lp) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for x in csv.DictReader(io.StringIO("".join(rows))):
        agg[short(x["Kernel Name"])][0] += 1
        agg[short(x["Kernel Name"])][1] += float(x["Metric Value"].replace(",", "")) * UNIT.get(x.get("Metric Unit", "nsecond"), 1e-3)
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(P, "r02_launches_c2.txt"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none  python bench.py --steps 5 --warmup 3 --no-extra --no-cpu-baseline\n")
        f.write("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes.  The timed region of that command is the\n"
                "# csmain_accumulate launches (one per step, resolve fused); the other kernels belong to the parity re-render, the\n"
                "# literal-frame leg and the L2 flush fills (torch).\n")
        f.write(f"{'launches':>8} {'total_us':>10} {'avg_us':>9} {'share':>6}  kernel\n")
        for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{c:8d} {t:10.1f} {t / c:9.2f} {100 * t / tot:5.1f}%  {n}\n")

# ---- csmain_accumulate from the bench command ------------------------------------------------------------------------
ks = load("r02_c2_accumulate.ncu-rep")
if ks:
    write_kernels("r02_c2_accumulate_ncu.txt",
                  "# ncu --set full --clock-control none --import-source on -k regex:csmain_accumulate -s 3 -c 2\n"
                  "#     python bench.py --steps 5 --warmup 3 --no-extra --no-cpu-baseline   (1920x1080, 64 spp per launch, resolve fused)\n", ks)
    d = ks[-1]
    samples = 1920 * 1080 * 64
    traffic["csmain_accumulate"] = {
        "dram_read_bytes_per_launch": val(d, "dram__bytes_read.sum"), "dram_write_bytes_per_launch": val(d, "dram__bytes_write.sum"),
        "dram_bytes_per_launch": val(d, "dram__bytes_read.sum") + val(d, "dram__bytes_write.sum"),
        "l2_write_bytes_per_launch": 32.0 * val(d, "lts__t_sectors_srcunit_tex_op_write.sum"),
        "warp_inst_per_launch": val(d, "smsp__inst_executed.sum"), "warp_inst_per_sample": val(d, "smsp__inst_executed.sum") / samples,
        "thread_inst_per_sample": val(d, "smsp__inst_executed.sum") * val(d, "smsp__thread_inst_executed_per_inst_executed.ratio") / samples,
        "active_threads_per_inst": val(d, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        "issue_active_pct": val(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"), "source": "profiles/r02_c2_accumulate_ncu.txt"}

# ---- the other literal-scene kernels (experiments/prof_c2.py) ----------------------------------------------------------
ks = load("r02_c2_literal.ncu-rep")
if ks:
    write_kernels("r02_c2_literal_ncu.txt",
                  "# ncu --set full --clock-control none --import-source on -c 40  python experiments/prof_c2.py   (1920x1080)\n"
                  "# 4x csmain_accumulate with the fused resolve, 1x without, resolve_kernel (linear, sRGB+BGRA), 3x csmain_frame_vec4,\n"
                  "# 3x csmain_frames_persistent (200 frames each), present_kernel; plus the context's self-test / row-span kernels\n", ks)
    for d in ks:
        n = short(d["_name"])
        if "csmain_frames_persistent" in n:
            traffic["csmain_frames_persistent"] = {
                "dram_bytes_per_frame": (val(d, "dram__bytes_read.sum") + val(d, "dram__bytes_write.sum")) / 200.0,
                "l2_write_bytes_per_frame": 32.0 * val(d, "lts__t_sectors_srcunit_tex_op_write.sum") / 200.0,
                "us_per_frame_under_ncu": val(d, "gpu__time_duration.sum") / 200.0, "source": "profiles/r02_c2_literal_ncu.txt"}
        for key in ("resolve_kernel", "present_kernel", "csmain_frame_vec4"):
            if key in n:
                traffic[key] = {"dram_read_bytes_per_launch": val(d, "dram__bytes_read.sum"),
                                "dram_write_bytes_per_launch": val(d, "dram__bytes_write.sum"),
                                "l2_write_bytes_per_launch": 32.0 * val(d, "lts__t_sectors_srcunit_tex_op_write.sum"),
                                "us_under_ncu": val(d, "gpu__time_duration.sum"), "source": "profiles/r02_c2_literal_ncu.txt"}

# ---- bounce kernels ------------------------------------------------------------------------------------------------------
for wl, pw, pixels in (("c3", 4, 1920 * 1080), ("c4", 1, 3840 * 2160)):
    ks = load(f"r02_{wl}_bounce.ncu-rep")
    if not ks:
        continue
    write_kernels(f"r02_{wl}_bounce_ncu.txt",
                  f"# ncu --set full --clock-control none --import-source on -k regex:bounce_kernel -c 40  python experiments/prof_scene.py {wl} {pw}\n"
                  f"# launch order: chunk-schedule wave (warm-up), chunk-schedule wave, default wave (warm-up), default wave; one wave = {pw}\n"
                  "# sample(s) of every pixel.  bounce_kernel_chunk = the chunk schedule, bounce_kernel = the queue schedule (tail launch)\n", ks)
    d0 = [i for i, d in enumerate(ks) if "<1," in d["_name"] or "<(bool)1" in d["_name"]]     # depth-0 launches open a wave
    if len(d0) < 4:
        continue
    for tag, sel in (("", ks[d0[3]:]), (":chunk_schedule", ks[d0[1]:d0[2]])):
        inst = sum(val(d, "smsp__inst_executed.sum") for d in sel)
        thr = sum(val(d, "smsp__inst_executed.sum") * val(d, "smsp__thread_inst_executed_per_inst_executed.ratio") for d in sel)
        traffic[f"bounce_kernel:{wl}{tag}"] = {
            "dram_bytes_per_sample_wave": sum(val(d, "dram__bytes_read.sum") + val(d, "dram__bytes_write.sum") for d in sel) / pw,
            "dram_read_bytes_per_sample_wave": sum(val(d, "dram__bytes_read.sum") for d in sel) / pw,
            "warp_inst_per_sample": inst / (pixels * pw), "active_threads_per_inst": thr / inst,
            "per_launch": [{"kernel": short(d["_name"]), "us_under_ncu": val(d, "gpu__time_duration.sum"),
                            "warp_inst": val(d, "smsp__inst_executed.sum"),
                            "active_threads_per_inst": val(d, "smsp__thread_inst_executed_per_inst_executed.ratio"),
                            "issue_active_pct": val(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                            "l1_hit_pct": val(d, "l1tex__t_sector_hit_rate.pct")} for d in sel],
            "source": f"profiles/r02_{wl}_bounce_ncu.txt"}

# ---- fused peer merge ------------------------------------------------------------------------------------------------------
for G_ in (2, 8):
    ks = load(f"r02_peer_merge_g{G_}.ncu-rep") or (load("r02_peer_merge.ncu-rep") if G_ == 2 else None)
    if not ks:
        continue
    fn = "r02_peer_merge_ncu.txt" if G_ == 2 else f"r02_peer_merge_g{G_}_ncu.txt"
    write_kernels(fn, "# ncu --set full --metrics nvlrx__bytes.sum,nvltx__bytes.sum --clock-control none -k regex:peer_merge\n"
                      f"#     PROF_PEER_DIRECT=1 python experiments/prof_peer.py   (ONE process, {G_} GPUs, one context each; 1920x1080,\n"
                      "#     64 spp per rank and frame; the merge kernel driven directly: ncu serialises launches, the rendezvous cannot run)\n", ks)
    d = ks[-1]
    traffic["peer_merge_kernel" if G_ == 2 else f"peer_merge_kernel:g{G_}"] = {
        "nvlink_rx_bytes_per_launch": val(d, "nvlrx__bytes.sum"), "nvlink_tx_bytes_per_launch": val(d, "nvltx__bytes.sum"),
        "dram_read_bytes_per_launch": val(d, "dram__bytes_read.sum"), "dram_write_bytes_per_launch": val(d, "dram__bytes_write.sum"),
        "us_under_ncu": val(d, "gpu__time_duration.sum"), "n_gpus": G_, "source": "profiles/" + fn}

json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
print("wrote", sorted(k for k in traffic if not k.startswith("_")))
