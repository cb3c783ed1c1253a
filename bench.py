#!/usr/bin/env python
"""bench.py — the measurement contract.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c2|c5|frame|c3|c4] [--no-extra]

Metric (BASELINE.json): Msamples/s (and ms/frame).  One sample = one evaluation of the per-pixel tracing path
(Test.hlsl CSMain, or one multi-bounce path for the scene workloads).  A "step" is one pass of the hot path over one frame.

  c2     BASELINE configs[1]: default scene, 1920x1080, 64 spp per step and GPU.  HEADLINE.  N=1: one launch per step
         (ptb_render_resolve: accumulate + resolve).  N>1: weak scaling in spp — every rank renders a disjoint 64-sample
         range of the frame per step, and EVERY step is merged by the pipelined fused peer-memory merge (ptb_peer_frame).
  c5     BASELINE configs[4]: 3840x2160, 1024 spp per step split over the N ranks (strong scaling), merged every step.
  c3     BASELINE configs[2]: 500 random spheres, 1920x1080, 256 spp per step, depth 8 (wavefront path tracer); N>1: sample
         ranges + peer merge (strong).
  c4     BASELINE configs[3]: 1,048,576 triangles, GPU-built BVH, 3840x2160, 256 spp per step; N>1: interleaved row bands
         per rank (tile sharding, strong), merge = NCCL reduce of disjoint RGBA8 rows.
  frame  the literal reference frame: 1 spp, no jitter, RGBA8 store (Renderer.ComputeFrame).

The default run measures c2 as the headline line and, in the SAME run, appends a `workloads` block with c5, c3 and c4
(fewer steps), each with its own roofline, parity check and — at N>1 — the one-GPU figure measured on rank 0 as the
strong-scaling basis.  `value` is device-timed (CUDA events on the launching stream, max over ranks); `e2e` goes through
the C ABI with HOST buffers (device->pinned-host read of the merged RGBA8 frame every step, wall clock, max over ranks).
`parity` re-renders what was merged on ONE GPU and compares bits (no oracle involved: the oracle-vs-GPU parity at these
sizes is tests/test_gpu_benched_configs.py).  `--impl reference` times the CPU restatement of the reference kernel
(oracle/, all host threads) on the same workload config.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FLOP_PER_SAMPLE = 31.0          # 27 (miss) + 0.1472 * 27 (hit extra), algorithmic
ACCUM_BYTES_PER_PIXEL = 32.0    # float4 read-modify-write per launch
RGBA8_BYTES_PER_PIXEL = 4.0
STATE_BYTES_PER_PATH = 40.0     # wavefront path state (2 x float4 + float2), read once and written once per segment
L2_BYTES = 126 * 1024 * 1024

WORKLOADS = {
    "c2": dict(width=1920, height=1080, spp=64, scaling="weak",
               name="BASELINE configs[1]: default scene 1920x1080, 64 spp/step"),
    "c5": dict(width=3840, height=2160, spp=1024, scaling="strong",
               name="BASELINE configs[4]: default scene 3840x2160, 1024 spp/step"),
    "frame": dict(width=1920, height=1080, spp=1, scaling="weak",
                  name="literal reference frame: 1920x1080, 1 spp, no jitter, RGBA8"),
    # scene workloads (SURVEY.md 8 a' rows n3-n7: no reference counterpart, oracle = this repo's own restatement)
    "c3": dict(width=1920, height=1080, spp=256, scene="spheres", max_depth=8, scaling="strong",
               name="BASELINE configs[2]: 500 random spheres (Lambert/metal/dielectric) 1920x1080, 256 spp/step, max depth 8"),
    "c4": dict(width=3840, height=2160, spp=256, scene="terrain", max_depth=8, scaling="strong",
               name="BASELINE configs[3]: 1,048,576-triangle terrain, GPU-built BVH, 3840x2160, 256 spp/step, max depth 8"),
}
DEFAULT_SHARD = {"c2": "p2p", "c5": "p2p", "c3": "p2p", "c4": "tiles", "frame": "samples"}
CPU_SAMPLES_PER_STEP_CAP = 1 << 30      # the CPU arms bound one step to ~1 Gi samples of the literal path


def make_scene(wl):
    from pathtracing_b200 import scenes
    if wl.get("scene") == "spheres":
        return scenes.random_spheres(500, seed=1, width=wl["width"], height=wl["height"])
    if wl.get("scene") == "terrain":
        return scenes.triangle_terrain(512, 2, seed=2, width=wl["width"], height=wl["height"])
    return None


def spp_per_step(name, wl, world):
    """Samples per pixel per STEP over all ranks: weak workloads grow with the world, strong ones are fixed."""
    return wl["spp"] * world if (wl["scaling"] == "weak" and name != "frame") else wl["spp"]


def workload_config(name, wl, world, shard):
    """What is measured — identical in the B200 line and in the `--impl reference` line (same keys, same values)."""
    if world == 1:
        sharding = "none"
    elif shard == "tiles":
        sharding = "interleaved bands of 4 image rows per rank; merge = NCCL reduce of disjoint RGBA8 rows, every step"
    elif shard == "p2p":
        sharding = ("disjoint sample ranges per rank; merge = ptb_peer_frame: one fused kernel per GPU over NVLink peer memory (P2P "
                    "loads of the partial accumulation buffers of the rank's pixel slab, rank-order sum, resolve, P2P store into "
                    "rank 0's image), rendezvous by peer-memory flag kernels, every step, pipelined behind the next step's render")
    else:
        sharding = "disjoint sample ranges per rank; merge = NCCL reduce(sum) of the float4 accumulation buffers + resolve on rank 0, every step"
    return {"workload": wl["name"], "width": wl["width"], "height": wl["height"],
            "spp_per_step_total": spp_per_step(name, wl, world), "seed": 0, "jitter": name != "frame",
            "max_depth": wl.get("max_depth", 0), "scene": wl.get("scene", "default (Test.hlsl literals)"),
            "scaling": wl["scaling"], "n_gpus": world, "sharding": sharding}


def use_host_cores():
    """torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU legs run on every core this process may use."""
    from oracle import pt_oracle as O
    if "TORCHELASTIC_RUN_ID" in os.environ or int(os.environ.get("WORLD_SIZE", "1")) > 1:
        O.set_threads(len(os.sched_getaffinity(0)))


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "sm_max_mhz": float(p.get("sm_max_mhz", 1965.0)), "source": "measured"}
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


def load_profile_constants():
    """ncu-derived per-kernel constants (profiles/traffic.json, regenerated by profiles/summarize*.py): DRAM bytes per
    launch and, for the scene kernels, warp instructions per sample and the SIMT efficiency."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def prof_entry(prof, key, scalar_name):
    """profiles/traffic.json entry as a dict (round-1 files stored a bare DRAM byte count)."""
    v = prof.get(key)
    if isinstance(v, dict):
        return v
    return {scalar_name: v} if v is not None else {}


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region, polled through NVML (the same counters
    `nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.*` prints) every 10 ms."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index = index
        self.sm, self.bits, self.power = [], 0, []
        self.mx = None
        self._stop = None
        self._thr = None

    def start(self):
        try:
            import threading
            import pynvml
            pynvml.nvmlInit()
            h = None
            try:   # CUDA_VISIBLE_DEVICES-safe: resolve by the CUDA device's PCI bus id when possible
                import torch
                pr = torch.cuda.get_device_properties(self.index)
                h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0")
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            self._stop = threading.Event()

            def loop():
                while not self._stop.is_set():
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        self.bits |= int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                        self.power.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
                    except Exception:
                        pass
                    self._stop.wait(0.01)
            self._thr = threading.Thread(target=loop, daemon=True)
            self._thr.start()
        except Exception:
            self._thr = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": self.mx, "reasons": [], "samples": 0, "power_w_max": None}
        if self._thr is None:
            return out
        self._stop.set()
        self._thr.join(timeout=2)
        if self.sm:
            out.update(sm_mhz=statistics.median(self.sm), samples=len(self.sm))
        if self.power:
            out["power_w_max"] = max(self.power)
        out["reasons"] = sorted(n for b, n in self.REASONS.items() if self.bits & b)
        return out


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


# ---- CPU arms (oracle = the reference's tracing logic restated; the ONLY place bench.py executes oracle/) -------------
ORACLE_BUILD = "gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp (no -march=native), OpenMP over rows (the Parallel.For analogue)"


def cpu_literal(name, wl, world, steps, warmup, one_thread=True):
    """The CPU restatement of Test.hlsl CSMain on `steps` steps of the workload: per step the SAME samples per pixel the
    B200 arm renders (capped at ~1 Gi samples per step, said so when it bites), jittered accumulate + resolve to RGBA8
    (c2 / c5) or the literal 1 spp RGBA8 frame.  -> dict for the cpu_baseline key."""
    import numpy as np
    from oracle import pt_oracle as O
    W, H = wl["width"], wl["height"]
    jitter = name != "frame"
    spp_full = spp_per_step(name, wl, world)
    spp = max(1, min(spp_full, CPU_SAMPLES_PER_STEP_CAP // (W * H)))
    p = O.frame_params(W, H)
    acc = np.zeros((H, W, 4), np.float32)
    threads = O.max_threads()
    t_acc, t_res, first = [], [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        if jitter:
            O.accumulate(p, first, spp, 0, True, accum=acc)
            t1 = time.perf_counter()
            O.resolve(acc, first + spp)
        else:
            O.csmain_frame(p, want_f32=False, want_rgba8=True, want_hit=False)
            t1 = time.perf_counter()
        t2 = time.perf_counter()
        first += spp
        if i >= warmup:
            t_acc.append(t1 - t0)
            t_res.append(t2 - t1)
    per_step = [a + b for a, b in zip(t_acc, t_res)]
    n = W * H * spp
    rates = sorted(n / t / 1e6 for t in per_step)
    out = {"value": n * len(per_step) / sum(per_step) / 1e6, "unit": "Msamples/s", "cores": threads, "kind": "port",
           "median": statistics.median(rates), "min": rates[0], "max": rates[-1],
           "ms_per_step": 1e3 * sum(per_step) / len(per_step),
           "accumulate_ms": 1e3 * sum(t_acc) / len(t_acc), "resolve_ms": 1e3 * sum(t_res) / len(t_res),
           "sample": (f"{W}x{H}, {spp} of {spp_full} spp per step x {len(per_step)} steps ({warmup} warm-up), "
                      + ("jittered accumulate + resolve to RGBA8, both under OpenMP" if jitter else "csmain frame -> RGBA8")
                      + f", {threads} threads; {ORACLE_BUILD}; {cpu_model()}")}
    if one_thread:
        spp1 = max(1, min(spp, 16))
        a1 = np.zeros((H, W, 4), np.float32)
        t0 = time.perf_counter()
        if jitter:
            O.accumulate(p, 0, spp1, 0, True, accum=a1, nthreads=1)
        else:
            O.csmain_frame(p, want_f32=False, want_rgba8=True, want_hit=False, nthreads=1)
        dt = time.perf_counter() - t0
        out["one_thread"] = {"value": W * H * spp1 / dt / 1e6, "unit": "Msamples/s",
                             "sample": f"{spp1} spp, 1 step, accumulate only" if jitter else "1 frame"}
    return out


def cpu_scene(wl, scene, steps, warmup):
    """The scene oracle (own CPU BVH, OpenMP over rows) on 1 spp per step (stated; Msamples/s of a path tracer does not
    depend on how many samples a step carries)."""
    import numpy as np
    from oracle import pt_oracle as O
    from oracle import pt_scene_oracle as S
    W, H = wl["width"], wl["height"]
    so = S.SceneOracle(scene, W, H, max_depth=wl["max_depth"])
    acc = np.zeros((H, W, 4), np.float32)
    times, first = [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        so.accumulate(first, 1, seed=0, jitter=True, accum=acc)
        dt = time.perf_counter() - t0
        first += 1
        if i >= warmup:
            times.append(dt)
    so.close()
    threads = O.max_threads()
    return {"value": W * H * len(times) / sum(times) / 1e6, "unit": "Msamples/s", "cores": threads, "kind": "port",
            "ms_per_step": 1e3 * sum(times) / len(times),
            "sample": (f"{W}x{H}, 1 of {wl['spp']} spp per step x {len(times)} steps, scene oracle (multi-bounce paths over its own "
                       f"CPU BVH), {threads} threads; {ORACLE_BUILD}; {cpu_model()}")}


def run_reference(args):
    """`--impl reference`: the reference's tracing logic on the box's host cores.  The reference's own implementation is
    HLSL via Vulkan (cannot run: no graphics stack) hosted by C# (no dotnet), so this is the oracle port with all host
    threads, on the B200 arm's config."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    use_host_cores()
    world = max(1, args.gpus)
    name = args.workload
    wl = dict(WORKLOADS[name])
    if args.spp:
        wl["spp"] = args.spp
    shard = args.shard or DEFAULT_SHARD[name]

    def one(nm, w, steps, warmup):
        if w.get("scene"):
            return cpu_scene(w, make_scene(w), max(1, min(steps, 3)), min(warmup, 1))
        return cpu_literal(nm, w, world, steps, warmup, one_thread=False)

    cpu = one(name, wl, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": cpu["value"], "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": cpu["ms_per_step"], "higher_is_better": True,
        "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, wl, world, shard),
        "cpu_baseline": cpu,
        "e2e": {"value": cpu["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if name == "c2" and not args.no_extra:
        extra = {}
        for nm in ("c5", "c3", "c4"):
            w = dict(WORKLOADS[nm])
            c = one(nm, w, 2, 1)
            extra[nm] = {"config": workload_config(nm, w, world, DEFAULT_SHARD[nm]), "value": c["value"], "unit": "Msamples/s",
                         "cpu_baseline": c}
        line["workloads"] = extra
    print(json.dumps(line))
    return 0


# ---- the B200 arm ---------------------------------------------------------------------------------------------------------
class Ctx:
    pass


def sha(t):
    import hashlib
    return hashlib.sha256(t.contiguous().cpu().numpy().tobytes()).hexdigest()[:16]


def measure(cx, name, K, Wm, headline):
    """One workload at cx.world ranks.  -> dict (rank 0) / None."""
    import torch
    import torch.distributed as dist
    from pathtracing_b200 import NativeTracer
    from pathtracing_b200.dist import GpuEngine, ShardedRenderer, sample_range

    world, rank, local, dev, stream, peaks = cx.world, cx.rank, cx.local, cx.dev, cx.stream, cx.peaks
    wl = dict(WORKLOADS[name])
    if cx.args.spp and headline:
        wl["spp"] = cx.args.spp
    W, H = wl["width"], wl["height"]
    n_pix = W * H
    literal = name == "frame"
    scene = make_scene(wl)
    depth = wl.get("max_depth", 0)
    spp_step = spp_per_step(name, wl, world)
    shard = (cx.args.shard if headline and cx.args.shard else DEFAULT_SHARD[name]) if world > 1 else "samples"
    p2p = world > 1 and shard == "p2p"
    fused = world == 1 and scene is None and not literal and not os.environ.get("PTB200_NO_FUSED_RESOLVE")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_img = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]   # double-buffered
    with torch.cuda.stream(stream):
        eng = GpuEngine(W, H, local, seed=0, stream=stream, scene=scene, max_depth=depth, peer_memory=p2p)
        t = eng.tracer
        sr = ShardedRenderer(eng, mode=shard)

        def step(s):
            first = s * spp_step
            if literal:
                t.compute_frame()
            elif fused:
                t.render_resolve(first, spp_step)
            elif p2p:
                sr.frame(first, spp_step, pipelined=True)
            else:
                sr.render(first, spp_step)
                sr.merge(0)

        # the pipelined merges are timed as ONE interval (nothing a merge does may hide in an untimed gap), L2-flush fills
        # included — unless the step's own buffers exceed L2 (two 16 B/pixel partials at 4K: 265 MB)
        pipelined = p2p
        big = 2 * n_pix * 16 > L2_BYTES
        flush = cx.flush if not pipelined else cx.flush[:160 * 1024 * 1024]
        do_flush = not (pipelined and big)

        # ---- device-timed value -------------------------------------------------------------------------------------
        for s in range(Wm):
            step(s)
        if p2p:
            sr.finish()
        fill_ms = 0.0
        if pipelined and do_flush:
            # the fills sit inside the one timed interval of a pipelined run; their own duration (alone on the GPU, median
            # of 5) is taken off again, so that N > 1 and N = 1 (fills untimed, between the per-step intervals) measure the
            # same thing.  What a merge adds while it shares the GPU with a fill stays in the interval.
            fe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
            torch.cuda.synchronize()
            for a, b in fe:
                a.record(stream)
                flush.fill_(7)
                b.record(stream)
            torch.cuda.synchronize()
            fill_ms = sorted(a.elapsed_time(b) for a, b in fe)[2]
        barrier()
        clocks = ClockSampler(local)
        if rank == 0 and headline:
            clocks.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K + 1)]
        launches0 = t.launch_count
        barrier()
        for s in range(K):
            if do_flush:
                flush.fill_(s & 0xFF)                  # L2 flush between timed iterations
            ev[s][0].record(stream)
            step(Wm + s)
            ev[s][1].record(stream)
        ev[K][0].record(stream)
        if p2p:
            sr.finish()                                # timed: drain the merges still in flight
        ev[K][1].record(stream)
        barrier()
        launches = t.launch_count - launches0
        step_ms = [a.elapsed_time(b) for a, b in ev[:K]]
        if pipelined:
            total_ms = ev[0][0].elapsed_time(ev[K][1]) - K * fill_ms
            kern_ms = step_ms                          # render launches only (the merge runs on its own stream)
        else:
            total_ms = sum(step_ms)
            kern_ms = step_ms
        clk = clocks.stop() if (rank == 0 and headline) else None

        # launch time of the dominant kernel alone (N>1 and unfused steps: the render part of a step)
        if world > 1 or not (fused or literal):
            kr = []
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for i in range(min(K, 3)):
                if p2p:
                    sr.finish()
                torch.cuda.synchronize()
                a.record(stream)
                st_, n_ = sample_range((Wm + K + i) * spp_step, spp_step, rank, world) if shard != "tiles" else ((Wm + K + i) * spp_step, spp_step)
                if p2p:
                    t.peer_frame(st_, n_, 0, render_only=True)
                else:
                    t.render_samples(st_, n_)
                b.record(stream)
                torch.cuda.synchronize()
                kr.append(a.elapsed_time(b))
                sr.spp_total += spp_step                  # merge what was just rendered (keeps the totals consistent)
                if p2p:
                    t.peer_frame(0, 0, sr.spp_total)
                else:
                    sr.merge(0)
            kern_ms = kr
            extra_steps = len(kr)
        else:
            extra_steps = 0

        # ---- e2e: C ABI with host buffers, image delivered to pinned host memory EVERY step -------------------------
        def step_e2e(s):
            step(s)
            if rank == 0:
                # queued on the library's copy stream behind the frame's LAST writer (the merge, at N>1): the PCIe transfer
                # overlaps the next step's kernels, and every copy completes inside the timed region
                t.read_rgba8_async(host_img[s & 1].data_ptr(), W * 4)
        e2e_steps = max(3, min(K, 50))
        base = Wm + K + extra_steps
        for s in range(Wm):
            step_e2e(base + s)
        if p2p:
            sr.finish()
        t.wait()
        barrier()
        t0 = time.perf_counter()
        for s in range(e2e_steps):
            step_e2e(base + Wm + s)
        if p2p:
            sr.finish()
        t.wait()                                       # last frame has landed in host memory
        barrier()
        e2e_s = time.perf_counter() - t0

        # ---- parity: what was merged vs a ONE-GPU re-render of the same sample sets ---------------------------------
        parity = None
        if not literal:
            parity = check_parity(cx, name, wl, scene, eng, sr, shard, spp_step)

        # ---- scene workloads: per-sample work counters from the counting build (untimed, 2 spp) ---------------------
        counters = None
        if scene is not None and rank == 0:
            with NativeTracer(W, H, device=local, seed=0, max_depth=depth, stream=stream.cuda_stream) as tc:
                tc.upload_scene(scene, count_stats=True)
                tc.render_samples(0, 2)
                tc.wait()
                st = tc.scene_stats()
                info = tc.scene_info()
            pth = float(st["paths"])
            counters = {"segments_per_sample": st["rays"] / pth, "nodes_visited_per_sample": st["nodes_visited"] / pth,
                        "prims_tested_per_sample": st["prims_tested"] / pth, "scatters_per_sample": st["scatters"] / pth,
                        "max_stack": st["max_stack"], "bvh_build_ms": info["build_ms"],
                        "nodes": info["sphere_nodes"] + info["triangle_nodes"],
                        "leaves": info["sphere_leaves"] + info["triangle_leaves"], "scene_bytes": info["device_bytes"]}

        # ---- strong scaling basis: the same workload on rank 0 alone, same run ---------------------------------------
        one_gpu = None
        if world > 1 and wl["scaling"] == "strong":
            barrier()
            if rank == 0:
                with NativeTracer(W, H, device=local, seed=0, max_depth=depth, stream=stream.cuda_stream) as t1:
                    if scene is not None:
                        t1.upload_scene(scene)
                    n1 = 2 if scene is not None else max(2, min(K, 5))
                    f1 = scene is None

                    def step1(s):
                        if f1:
                            t1.render_resolve(s * spp_step, spp_step)
                        else:
                            t1.render_samples(s * spp_step, spp_step)
                            t1.resolve()
                    step1(0)
                    torch.cuda.synchronize()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(stream)
                    for s in range(n1):
                        step1(1 + s)
                    b.record(stream)
                    torch.cuda.synchronize()
                    ms1 = a.elapsed_time(b) / n1
                one_gpu = {"value": n_pix * spp_step / (ms1 * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_frame": ms1, "steps": n1,
                           "note": "the same frame on rank 0 alone, measured in this run while the other ranks idle"}
            barrier()

        kmode = t.kernel_mode
        eng.close()

    tot = torch.tensor([total_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(tot[0]), float(tot[1])
    if rank != 0:
        return None

    samples_per_step = n_pix * spp_step                  # all ranks together
    value = samples_per_step * K / (total_ms * 1e-3) / 1e6
    e2e_value = samples_per_step * e2e_steps / (e2e_ms * 1e-3) / 1e6
    spp_rank = spp_step / world if shard != "tiles" else spp_step   # per-pixel samples one rank's launch evaluates
    pix_rank = n_pix if shard != "tiles" else n_pix / world
    k_ms = statistics.mean(kern_ms)
    prof = cx.prof

    # ---- roofline of the dominant kernel, from the live launch durations ---------------------------------------------
    if literal:
        alg = RGBA8_BYTES_PER_PIXEL * n_pix
        achieved = alg / (k_ms * 1e-3) / 1e9
        roof = {"kernel": "csmain_frame_vec4", "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"],
                "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": prof_entry(prof, "csmain_frame_vec4", "dram_bytes_per_launch").get("dram_bytes_per_launch"),
                "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']})",
                "algorithmic_bytes_per_launch": alg, "launch_ms": k_ms}
    elif scene is not None:
        c = counters
        prim_b = 16.0 if wl["scene"] == "spheres" else 48.0
        b_sample = (c["nodes_visited_per_sample"] * 64.0 + c["prims_tested_per_sample"] * prim_b +
                    c["segments_per_sample"] * 2.0 * STATE_BYTES_PER_PATH + ACCUM_BYTES_PER_PIXEL)
        wave_ms = k_ms / max(1.0, spp_rank)              # time per sample of every pixel of the rank
        samples_per_s = pix_rank / (wave_ms * 1e-3)
        stream_b = c["segments_per_sample"] * 2.0 * STATE_BYTES_PER_PATH + 32.0 + ACCUM_BYTES_PER_PIXEL / max(1.0, min(64.0, spp_rank))
        stream_gbs = stream_b * samples_per_s / 1e9
        pk = prof_entry(prof, f"bounce_kernel:{name}", "dram_bytes_per_sample_wave")
        issue_peak = 148 * 4 * peaks["sm_max_mhz"] * 1e6 / 1e9          # G warp-instructions / s
        inst = pk.get("warp_inst_per_sample")
        simt = pk.get("active_threads_per_inst")
        roof = {"kernel": "bounce_kernel", "bound": "issue slots x SIMT efficiency (divergent traversal; the scene is L2/L1 resident)",
                "unit": "G warp-inst/s", "peak": issue_peak,
                "achieved": inst * samples_per_s / 1e9 if inst else None,
                "frac": inst * samples_per_s / 1e9 / issue_peak if inst else None,
                "simt_efficiency": simt / 32.0 if simt else None,
                "useful_lane_frac": (inst * samples_per_s / 1e9 / issue_peak) * (simt / 32.0) if inst and simt else None,
                "traffic": pk.get("dram_bytes_per_sample_wave"),
                "peak_source": f"148 SM x 4 schedulers x sm_max_mhz from MEASURED_PEAKS.json ({peaks['source']}); achieved = warp "
                               "instructions per sample (ncu smsp__inst_executed.sum of one whole wave / its samples, profiles/traffic.json) "
                               "x live samples/s; simt_efficiency = ncu smsp__thread_inst_executed_per_inst_executed / 32",
                "algorithmic_bytes_per_sample": b_sample, "algorithmic_gbs": b_sample * samples_per_s / 1e9,
                "hbm_stream_bytes_per_sample": stream_b, "hbm_stream_gbs": stream_gbs, "hbm_stream_frac": stream_gbs / peaks["hbm_gbs"],
                "hbm_peak": peaks["hbm_gbs"], "launches_per_step": max(1.0, launches / K), "wave_ms": wave_ms,
                "note": "algorithmic bytes (64 B per node visit, 16/48 B per primitive test, 2 x 40 B path state per segment, 32 B "
                        "accumulation) are mostly L1/L2 hits; hbm_stream_* is the share that must come from HBM",
                "counters": c}
    else:
        fp32_peak = 148 * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12     # TFLOP/s at max SM clock
        flops = FLOP_PER_SAMPLE * pix_rank * spp_rank
        achieved = flops / (k_ms * 1e-3) / 1e12
        pk = prof_entry(prof, "csmain_accumulate", "dram_bytes_per_launch")
        inst = pk.get("warp_inst_per_sample")
        issue_peak = 148 * 4 * peaks["sm_max_mhz"] * 1e6 / 1e9
        sps = pix_rank * spp_rank / (k_ms * 1e-3)
        roof = {"kernel": "csmain_accumulate", "bound": "fp32", "achieved": achieved, "peak": fp32_peak,
                "unit": "TFLOP/s", "frac": achieved / fp32_peak,
                "traffic": pk.get("dram_bytes_per_launch"),
                "peak_source": f"148 SM x 128 lanes x 2 x sm_max_mhz from MEASURED_PEAKS.json ({peaks['source']}); "
                               "FP32 CUDA-core bound: no tensor or HBM roof applies (SURVEY.md 8d, DESIGN.md 4)",
                "algorithmic_flops_per_launch": flops, "launch_ms": k_ms,
                "issue_slot_frac": inst * sps / 1e9 / issue_peak if inst else None,
                "issue_note": "issue_slot_frac = ncu warp instructions per sample (profiles/traffic.json) x live samples/s / "
                              "(148 SM x 4 schedulers x sm_max_mhz): the exact IEEE sequences cost more instructions than the 31 "
                              "algorithmic flop, so this is the fairer utilisation figure",
                "hbm_bytes_per_launch": ACCUM_BYTES_PER_PIXEL * pix_rank,
                "hbm_frac": ACCUM_BYTES_PER_PIXEL * pix_rank / (k_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}

    cfg = workload_config(name, wl, world, shard)
    cfg.update({
        "spp_per_step_per_rank": spp_rank,
        "step": ("ptb_render_resolve: accumulate + resolve in one launch" if fused else "ptb_compute_frame" if literal else
                 "ptb_peer_frame (render + pipelined fused peer merge)" if p2p else "ptb_render_samples + merge / ptb_resolve"),
        "l2": ("no flush: each step streams two 16 B/pixel partial buffers (> 126 MB L2) plus the peers' slabs" if (pipelined and big) else
               "flushed between timed steps (" + (f"160 MB fill inside the timed interval, its own {fill_ms * 1e3:.1f} us per step taken off again)"
                                                  if pipelined else "256 MB fill, untimed)")),
        "timing": ("one CUDA-event interval over all K steps and the drain of in-flight merges, minus K x the isolated duration of the "
                   "L2-flush fill; max over ranks" if pipelined else "per-step CUDA events on the launching stream, summed; max over ranks")})
    res = {"config": cfg, "value": value, "unit": "Msamples/s", "ms_per_step": total_ms / K, "ms_per_frame": total_ms / K,
           "steps": K, "warmup": Wm, "scaling": wl["scaling"], "roofline": roof,
           "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": 128, "d2h_bytes_per_step": n_pix * 4,
                   "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                   "note": "C-ABI calls with host buffers; per step the merged RGBA8 frame is read into double-buffered pinned host "
                           "memory (ptb_read_rgba8_async, overlapping the next step) and the last copy completes inside the timed "
                           "region.  The path's only input is the parameter block (Test.hlsl reads nothing)"},
           "gpu_launches": int(launches), "kernel_mode": kmode, "parity": parity, "clocks": clk}
    if one_gpu:
        res["one_gpu"] = one_gpu
        res["speedup_vs_one_gpu"] = value / one_gpu["value"]
    return res


def check_parity(cx, name, wl, scene, eng, sr, shard, spp_step):
    """Re-render on rank 0 alone what the N ranks rendered and merged, and compare bits.
      N = 1   the benched path (fused launch / batched waves) vs the plain two-call path on a fresh context;
      p2p     merged running total == per-rank partials added in rank order, frame after frame (float bits), and the RGBA8
              image delivered to the host by the LAST e2e step == its resolve;
      tiles   merged RGBA8 == a one-GPU render of the whole frame (identical bytes);
      samples NCCL reduce: identical for N = 2, else <= 1e-6 relative (the reduction order is NCCL's)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from pathtracing_b200 import NativeTracer
    from pathtracing_b200.dist import sample_range
    world, rank, local, stream = cx.world, cx.rank, cx.local, cx.stream
    W, H = wl["width"], wl["height"]
    depth = wl.get("max_depth", 0)
    t = eng.tracer
    out = {"checked": True, "ok": None, "mismatched_pixels": None, "what": None}
    # a fresh, short sequence through exactly the benched code path: 2 frames of the workload's spp per step
    frames = 2 if scene is None else 1
    sr.engine.accum_reset()
    sr.spp_total = 0
    if world > 1:
        dist.barrier()
    fused = world == 1 and scene is None
    for f in range(frames):
        if fused:
            t.render_resolve(f * spp_step, spp_step)
        elif shard == "p2p" and world > 1:
            sr.frame(f * spp_step, spp_step, pipelined=True)
        else:
            sr.render(f * spp_step, spp_step)
            sr.merge(0)
    if shard == "p2p" and world > 1:
        sr.finish()
    t.wait()
    got_img = t.read_rgba8() if rank == 0 else None
    slabs = None
    if shard == "p2p" and world > 1:
        mine = t.peer_read_total()
        slabs = [None] * world
        dist.all_gather_object(slabs, mine)
    elif rank == 0:
        got_acc = t.read_accum() if world == 1 else eng.accum_tensor().cpu().numpy()
    if rank == 0:
        with NativeTracer(W, H, device=local, seed=0, max_depth=depth, stream=stream.cuda_stream) as r1:
            if scene is not None:
                r1.upload_scene(scene)
                if world == 1:
                    r1.set_wave_samples(1)                  # one sample per wave: the unbatched schedule
            total = np.zeros((H, W, 4), np.float32)
            for f in range(frames):
                if world > 1 and shard in ("p2p", "samples"):
                    part_sum = None
                    for r in range(world):
                        st_, n_ = sample_range(f * spp_step, spp_step, r, world)
                        r1.accum_reset()
                        if n_:
                            r1.render_samples(st_, n_)
                        pr = r1.read_accum()
                        part_sum = pr if part_sum is None else part_sum + pr          # rank order
                    total = total + part_sum
                else:
                    r1.accum_reset()
                    r1.render_samples(f * spp_step, spp_step)
                    total = total + r1.read_accum() if frames > 1 or world > 1 else r1.read_accum()
            r1.write_accum(total, frames * spp_step)
            r1.resolve()
            want_img = r1.read_rgba8()
        bad_img = int(np.count_nonzero((got_img != want_img).any(axis=-1)))
        if slabs is not None:
            got = np.zeros((H * W, 4), np.float32)
            for first, s_ in slabs:
                got[first:first + len(s_)] = s_
            got_acc = got.reshape(H, W, 4)
        if shard == "tiles" and world > 1:
            bad_acc = None                                   # tiles merge images, not accumulation buffers
            ok = bad_img == 0
            what = "merged RGBA8 image == one-GPU render of the whole frame (bytes)"
        elif shard == "samples" and world > 2:
            rel = np.abs(got_acc - total) / np.maximum(np.abs(total), 1e-6)
            bad_acc = int(np.count_nonzero((rel > 1e-6).any(axis=-1)))
            ok = bad_acc == 0 and int(np.abs(got_img.astype(int) - want_img.astype(int)).max()) <= 1
            what = "NCCL reduce: accumulation within 1e-6 relative of the rank-ordered sum, RGBA8 within 1 code"
        else:
            bad_acc = int(np.count_nonzero((got_acc.view(np.uint32) != total.view(np.uint32)).any(axis=-1)))
            ok = bad_img == 0 and bad_acc == 0
            what = ("fused launch / batched waves == plain per-call path (float bits + RGBA8 bytes)" if world == 1 else
                    "merged accumulation == per-rank partials added in rank order on one GPU (float bits); merged RGBA8 == its resolve")
        out.update(ok=bool(ok), mismatched_pixels=bad_img if bad_acc is None else max(bad_img, bad_acc), what=what,
                   frames=frames, spp_per_frame=spp_step, image_sha256_16=sha(torch.from_numpy(got_img)))
    if world > 1:
        dist.barrier()
    return out if rank == 0 else None


def literal_frame_block(cx):
    """The literal reference frame (1 spp, RGBA8) as a secondary block of the c2 line."""
    import torch
    from pathtracing_b200 import NativeTracer
    W, H = 1920, 1080
    n_pix = W * H
    stream, peaks = cx.stream, cx.peaks
    with torch.cuda.stream(stream), NativeTracer(W, H, device=cx.local, stream=stream.cuda_stream) as tf:
        for _ in range(5):
            tf.compute_frame()
        tf.compute_frames(5)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_bb = 200
        a.record(stream)
        for _ in range(n_bb):
            tf.compute_frame()
        b.record(stream)
        torch.cuda.synchronize()
        bb_ms = a.elapsed_time(b) / n_bb
        a.record(stream)
        tf.compute_frames(n_bb)
        b.record(stream)
        torch.cuda.synchronize()
        per_ms = a.elapsed_time(b) / n_bb
        iso = []
        for i in range(20):
            cx.flush.fill_(i)
            a.record(stream)
            tf.compute_frame()
            b.record(stream)
            torch.cuda.synchronize()
            iso.append(a.elapsed_time(b))
        iso_ms = statistics.median(iso)
        bytes_frame = RGBA8_BYTES_PER_PIXEL * n_pix
        return {"workload": WORKLOADS["frame"]["name"],
                "ms_per_frame_back_to_back": bb_ms, "ms_per_frame_persistent_launch": per_ms, "ms_per_frame_l2_flushed": iso_ms,
                "msamples_per_s_back_to_back": n_pix / (bb_ms * 1e-3) / 1e6, "msamples_per_s_persistent_launch": n_pix / (per_ms * 1e-3) / 1e6,
                "roofline": {"kernel": "csmain_frames_persistent", "bound": "hbm", "unit": "GB/s",
                             "achieved": bytes_frame / (per_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"],
                             "frac": bytes_frame / (per_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes_per_frame": bytes_frame,
                             "traffic": prof_entry(cx.prof, "csmain_frames_persistent", "dram_bytes_per_frame").get("dram_bytes_per_frame"),
                             "note": f"ptb_compute_frames({n_bb}): {n_bb} frames in one persistent launch (every frame computed and stored in "
                                     "full); ptb_compute_frame back to back is one ~3 us kernel per launch and stays launch-latency bound"}}


def run_b200(args):
    import torch
    import torch.distributed as dist

    cx = Ctx()
    cx.args = args
    cx.world = int(os.environ.get("WORLD_SIZE", "1"))
    cx.rank = int(os.environ.get("RANK", "0"))
    cx.local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    torch.cuda.set_device(cx.local)
    cx.dev = torch.device("cuda", cx.local)
    if cx.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=cx.dev)
    cx.peaks = load_peaks()
    cx.prof = load_profile_constants()
    cx.stream = torch.cuda.Stream(device=cx.dev)
    cx.flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=cx.dev)                   # > 126 MB L2
    name = args.workload
    K, Wm = args.steps, args.warmup

    head = measure(cx, name, K, Wm, headline=True)
    extra = {}
    if name == "c2" and not args.no_extra:
        for nm, k_, w_ in (("c5", max(3, min(K, 10)), 3), ("c3", max(2, min(K, 3)), 1), ("c4", 2, 1)):
            extra[nm] = measure(cx, nm, k_, w_, headline=False)
    lit = None
    if cx.rank == 0 and cx.world == 1 and name == "c2":
        lit = literal_frame_block(cx)

    if cx.rank == 0:
        wl = WORKLOADS[name]
        cpu = None
        if cx.world == 1 and not args.no_cpu_baseline:
            use_host_cores()
            cpu = (cpu_scene(wl, make_scene(wl), 2, 1) if wl.get("scene") else
                   cpu_literal(name, dict(wl, spp=args.spp or wl["spp"]), 1, 10, 2))
            if name == "c2" and not args.no_extra:
                for nm in extra:
                    w = WORKLOADS[nm]
                    extra[nm]["cpu_baseline"] = (cpu_scene(w, make_scene(w), 1, 1) if w.get("scene") else cpu_literal(nm, w, 1, 2, 1, False))
        line = {
            "metric": "Msamples/s", "value": head["value"], "unit": "Msamples/s", "n_gpus": cx.world, "steps": K, "warmup": Wm,
            "ms_per_step": head["ms_per_step"], "ms_per_frame": head["ms_per_frame"], "higher_is_better": True,
            "scaling": head["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": head["config"], "roofline": head["roofline"], "cpu_baseline": cpu, "e2e": head["e2e"],
            "gpu_launches": head["gpu_launches"], "kernel_mode": head["kernel_mode"], "clocks": head["clocks"],
            "parity": head["parity"],
            "vs_cpu_baseline": ({"device": head["value"] / cpu["value"], "e2e": head["e2e"]["value"] / cpu["value"],
                                 "note": "this run's own GPU figures / this run's cpu_baseline.value (same config, same box); "
                                         "vs_baseline stays null: the reference publishes no number"} if cpu else None),
            "workloads": {k: {kk: vv for kk, vv in v.items() if kk != "clocks"} for k, v in extra.items() if v} or None,
            "literal_frame": lit,
        }
        for k in ("one_gpu", "speedup_vs_one_gpu"):
            if k in head:
                line[k] = head[k]
        print(json.dumps(line))
    if cx.world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel per step")
    ap.add_argument("--cpu-spp", type=int, default=0, help="(ignored; kept for old command lines: the CPU arms run the workload's own spp)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="c2 only: skip the workloads block (c5, c3, c4)")
    ap.add_argument("--shard", default=None, choices=["samples", "tiles", "p2p"],
                    help="multi-GPU partitioning of the headline workload: sample ranges merged by the pipelined fused peer-memory "
                         "kernel (p2p, default for c2/c5/c3), interleaved image row bands (tiles, default for c4), or sample ranges "
                         "merged by an NCCL reduce (samples)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3           # timing rule: W >= 3
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29513", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
