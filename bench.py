#!/usr/bin/env python
"""bench.py — the measurement contract.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c2|c5|frame]

Metric (BASELINE.json): Msamples/s (and ms/frame).  One sample = one evaluation of the per-pixel
tracing path (Test.hlsl CSMain).  Workloads:

  c2     BASELINE configs[1]: default scene, 1920x1080, 64 spp per step (jittered samples summed
         in-kernel into the float4 accumulation buffer, then resolved to RGBA8).  DEFAULT at N=1.
         At N>1 every rank renders a disjoint 64-sample range of the same frame per step (weak
         scaling in spp); the per-rank accumulation buffers are merged with one NCCL reduce + one
         resolve at the end of the K steps, INSIDE the timed total.
  c5     BASELINE configs[4]: 3840x2160, 1024 spp per step split across the N ranks (strong
         scaling), NCCL reduce + resolve every step.
  frame  the literal reference frame: 1 spp, no jitter, RGBA8 store (Renderer.ComputeFrame).

A "step" is one pass of the hot path over one frame.  `value` is device-timed (CUDA events on the
launching stream, L2 flushed between timed steps, max over ranks); `e2e` goes through the C ABI
with HOST buffers (device->pinned-host read of the image every step, wall clock, max over ranks).
`--impl reference` times the CPU restatement of the reference kernel (oracle/, all host threads)
on a bounded sample of the same workload.
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

WORKLOADS = {
    "c2": dict(width=1920, height=1080, spp=64, name="BASELINE configs[1]: default scene 1920x1080, 64 spp/step"),
    "c5": dict(width=3840, height=2160, spp=1024, name="BASELINE configs[4]: default scene 3840x2160, 1024 spp/step"),
    "frame": dict(width=1920, height=1080, spp=1, name="literal reference frame: 1920x1080, 1 spp, no jitter, RGBA8"),
    # scene workloads (SURVEY.md 8 a' rows n3-n7: no reference counterpart, oracle = this repo's own restatement)
    "c3": dict(width=1920, height=1080, spp=256, scene="spheres", max_depth=8,
               name="BASELINE configs[2]: 500 random spheres (Lambert/metal/dielectric) 1920x1080, 256 spp/step, max depth 8"),
    "c4": dict(width=3840, height=2160, spp=256, scene="terrain", max_depth=8,
               name="BASELINE configs[3]: 1,048,576-triangle terrain, GPU-built BVH, 3840x2160, 256 spp/step, max depth 8"),
}
STATE_BYTES_PER_PATH = 40.0     # wavefront path state (2 x float4 + float2), read once and written once per segment


def make_scene(wl):
    from pathtracing_b200 import scenes
    if wl.get("scene") == "spheres":
        return scenes.random_spheres(500, seed=1, width=wl["width"], height=wl["height"])
    if wl.get("scene") == "terrain":
        return scenes.triangle_terrain(512, 2, seed=2, width=wl["width"], height=wl["height"])
    return None


def cpu_scene_sample(wl, scene, spp_sample, steps, warmup):
    """Times the scene oracle (own CPU BVH, OpenMP over rows) on `spp_sample` samples per pixel per step."""
    import numpy as np
    from oracle import pt_oracle as O
    from oracle import pt_scene_oracle as S
    W, H = wl["width"], wl["height"]
    so = S.SceneOracle(scene, W, H, max_depth=wl["max_depth"])
    acc = np.zeros((H, W, 4), np.float32)
    times, first = [], 0
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        so.accumulate(first, spp_sample, seed=0, jitter=True, accum=acc)
        dt = time.perf_counter() - t0
        first += spp_sample
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    so.close()
    return W * H * spp_sample * len(times) / total / 1e6, 1e3 * total / len(times), O.max_threads()


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
            # CUDA_VISIBLE_DEVICES-safe: resolve by the CUDA device's UUID/PCI bus id when possible
            h = None
            try:
                import torch
                bus = torch.cuda.get_device_properties(self.index).pci_bus_id
                dom = torch.cuda.get_device_properties(self.index).pci_domain_id
                dev = torch.cuda.get_device_properties(self.index).pci_device_id
                h = pynvml.nvmlDeviceGetHandleByPciBusId(f"{dom:08x}:{bus:02x}:{dev:02x}.0")
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


# ------------------------------------------------------------------------------------------------
def cpu_sample(width, height, spp_sample, steps, warmup, jitter=True):
    """Times the CPU oracle (the C restatement of Test.hlsl CSMain) with all host threads on
    `spp_sample` samples per pixel of the workload's frame per step.  -> (Msamples/s, ms/step, threads)"""
    import numpy as np
    from oracle import pt_oracle as O
    p = O.frame_params(width, height)
    acc = np.zeros((height, width, 4), np.float32)
    threads = O.max_threads()
    first = 0
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        if jitter:
            O.accumulate(p, first, spp_sample, 0, True, accum=acc)
            O.resolve(acc, first + spp_sample)
        else:
            O.csmain_frame(p, want_f32=False, want_rgba8=True, want_hit=False)
        dt = time.perf_counter() - t0
        first += spp_sample
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    msamples = width * height * spp_sample * len(times) / total / 1e6
    return msamples, 1e3 * total / len(times), threads


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def run_reference(args):
    """`--impl reference`: the reference's tracing logic on the box's host cores.  The reference's own
    implementation is HLSL via Vulkan (cannot run: no graphics stack) hosted by C# (no dotnet), so this
    is the oracle port, OpenMP over rows (the Parallel.For analogue), all threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    use_host_cores()
    wl = dict(WORKLOADS[args.workload])
    if args.spp:
        wl["spp"] = args.spp
    W, H, spp = wl["width"], wl["height"], wl["spp"]
    jitter = args.workload != "frame"
    if wl.get("scene"):
        spp_sample = 1
        ms, ms_step, threads = cpu_scene_sample(wl, make_scene(wl), spp_sample, min(args.steps, 3), min(args.warmup, 1))
        sample = (f"{W}x{H}, {spp_sample} of {spp} spp per step (scene oracle: multi-bounce paths over its own CPU BVH), "
                  f"{min(args.steps, 3)} steps; Msamples/s is spp-invariant")
    else:
        spp_sample = 1 if not jitter else max(1, min(spp, args.cpu_spp))
        ms, ms_step, threads = cpu_sample(W, H, spp_sample, args.steps, args.warmup, jitter)
        sample = (f"{W}x{H}, {spp_sample} of {spp} spp per step ({'jittered accumulate + resolve' if jitter else 'csmain frame'}), "
                  f"{args.steps} steps; Msamples/s is spp-invariant")
    line = {
        "impl": "reference", "metric": "Msamples/s", "value": ms, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "cpu_model": cpu_model()},
        "cpu_baseline": {"value": ms, "unit": "Msamples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": ms, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from pathtracing_b200 import NativeTracer
    from pathtracing_b200.dist import GpuEngine, ShardedRenderer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        opts = None
        try:   # NCCL kernels on a high-priority stream: the rendezvous of a pipelined merge must not queue behind render blocks
            opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        except Exception:
            opts = None
        dist.init_process_group("nccl", device_id=dev, pg_options=opts)

    wl = dict(WORKLOADS[args.workload])
    if args.spp:
        wl["spp"] = args.spp
    W, H, spp = wl["width"], wl["height"], wl["spp"]
    scene = make_scene(wl)
    n_pix = W * H
    strong = args.workload in ("c5", "c4")     # c4 (BASELINE configs[3]) is ONE 256-spp 4K frame split over the GPUs
    literal = args.workload == "frame"
    # samples per pixel per STEP over all ranks: c2 scales with the world (weak), c5 is fixed (strong)
    spp_step = spp if (strong or literal) else spp * world
    K, Wm = args.steps, args.warmup
    peaks = load_peaks()

    stream = torch.cuda.Stream(device=dev)
    host_img = [torch.empty((H, W, 4), dtype=torch.uint8).pin_memory() for _ in range(2)]   # double-buffered
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)                   # > 126 MB L2

    with torch.cuda.stream(stream):
        eng = GpuEngine(W, H, local, seed=0, stream=stream, scene=scene, max_depth=wl.get("max_depth", 0),
                        peer_memory=args.shard == "p2p")
        t = eng.tracer
        sr = ShardedRenderer(eng, mode=args.shard)
        merge_every_step = strong or world == 1        # c2 at N>1: one merge for the whole progressive render

        def barrier():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        pipelined = args.shard == "p2p" and not os.environ.get("PTB200_NO_PIPELINE")      # merge of frame k overlaps the rendering of frame k+1 (double-buffered partials)

        def do_merge():
            if pipelined:
                sr.merge(0, pipelined=True)
            else:
                sr.merge(0)

        # one GPU, literal scene: accumulate + resolve are ONE launch (ptb_render_resolve; PTB200_NO_FUSED_RESOLVE=1 for A/B)
        fused = world == 1 and scene is None and not literal and not os.environ.get("PTB200_NO_FUSED_RESOLVE")

        def step_device(step):
            if literal:
                t.compute_frame()
                return
            if fused:
                t.render_resolve(step * spp_step, spp_step)
                return
            sr.render(step * spp_step, spp_step)
            if merge_every_step:
                do_merge()

        # ---- device-timed value ---------------------------------------------------------------------
        for s in range(Wm):
            step_device(s)
        if not literal and not merge_every_step:
            do_merge()                                 # progressive merge of the warm-up samples: the timed merge below is not NCCL's first call
        if pipelined:
            sr.finish()
        barrier()
        clocks = ClockSampler(local)
        if rank == 0:
            clocks.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K + 1)]
        kern_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        launches0 = t.launch_count
        # pipelined merges are timed as ONE interval, so a flush fill would sit inside the timed region; when the
        # double-buffered partials alone exceed the 126 MB L2 every step already streams more than L2 and no flush is needed
        skip_flush = pipelined and 2 * n_pix * 16 > 126 * 1024 * 1024
        barrier()
        for s in range(K):
            if not skip_flush:
                flush.fill_(s & 0xFF)                  # L2 flush between timed iterations (untimed)
            ev[s][0].record(stream)
            kern_ev[s][0].record(stream)
            if literal:
                t.compute_frame()
            elif fused:
                t.render_resolve((Wm + s) * spp_step, spp_step)
            else:
                sr.render((Wm + s) * spp_step, spp_step)
            kern_ev[s][1].record(stream)
            if not literal and not fused and merge_every_step:
                do_merge()
            ev[s][1].record(stream)
        ev[K][0].record(stream)
        if not literal and not merge_every_step:
            sr.merge(0)                                # timed: NCCL reduce over NVLink + resolve on rank 0
        if pipelined:
            sr.finish()                                # timed: drain the merges still in flight
        ev[K][1].record(stream)
        barrier()
        launches = t.launch_count - launches0
        step_ms = [a.elapsed_time(b) for a, b in ev[:K]]
        merge_ms = ev[K][0].elapsed_time(ev[K][1])
        kern_ms = [a.elapsed_time(b) for a, b in kern_ev]
        total_ms = sum(step_ms) + merge_ms
        if pipelined:
            # merges overlap the following step (and the L2 flush between steps): time the whole loop as ONE interval,
            # flush fills included, so nothing a merge does can hide in an untimed gap
            total_ms = ev[0][0].elapsed_time(ev[K][1])
        clk = clocks.stop() if rank == 0 else None

        # ---- e2e: C ABI with host buffers, image delivered to pinned host memory EVERY step -------------
        def step_e2e(step):
            if literal:
                t.compute_frame()
            elif fused:
                t.render_resolve(step * spp_step, spp_step)
            else:
                sr.render(step * spp_step, spp_step)
                sr.merge(0)
            if rank == 0:
                # queued on the library's copy stream: the PCIe transfer overlaps the next step's kernels,
                # and every copy completes inside the timed region
                t.read_rgba8_async(host_img[step & 1].data_ptr(), W * 4)
        e2e_steps = max(3, min(K, 50))
        base = Wm + K
        for s in range(Wm):
            step_e2e(base + s)
        t.wait()
        barrier()
        t0 = time.perf_counter()
        for s in range(e2e_steps):
            step_e2e(base + Wm + s)
        t.wait()                                       # last frame has landed in host memory
        barrier()
        e2e_s = time.perf_counter() - t0

        # ---- scene workloads: per-sample work counters from the counting build (untimed, 2 spp) ----------------
        scene_counters = None
        if scene is not None and rank == 0:
            with NativeTracer(W, H, device=local, seed=0, max_depth=wl["max_depth"], stream=stream.cuda_stream) as tc:
                tc.upload_scene(scene, count_stats=True)
                tc.render_samples(0, 2)
                tc.wait()
                st = tc.scene_stats()
                info = tc.scene_info()
            p = float(st["paths"])
            scene_counters = {"segments_per_sample": st["rays"] / p, "nodes_visited_per_sample": st["nodes_visited"] / p,
                              "prims_tested_per_sample": st["prims_tested"] / p, "scatters_per_sample": st["scatters"] / p,
                              "max_stack": st["max_stack"], "bvh_build_ms": info["build_ms"],
                              "nodes": info["sphere_nodes"] + info["triangle_nodes"],
                              "leaves": info["sphere_leaves"] + info["triangle_leaves"], "scene_bytes": info["device_bytes"]}

        # ---- the literal reference frame (1 spp, RGBA8) as a secondary line ------------------------------
        literal_frame = None
        if rank == 0 and world == 1 and not literal and scene is None:
            with NativeTracer(W, H, device=local, stream=stream.cuda_stream) as tf:
                for _ in range(5):
                    tf.compute_frame()
                torch.cuda.synchronize()
                n_bb = 200
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                for _ in range(n_bb):
                    tf.compute_frame()
                b.record(stream)
                torch.cuda.synchronize()
                bb_ms = a.elapsed_time(b) / n_bb
                iso = []
                for i in range(20):
                    flush.fill_(i)
                    a.record(stream)
                    tf.compute_frame()
                    b.record(stream)
                    torch.cuda.synchronize()
                    iso.append(a.elapsed_time(b))
                iso_ms = statistics.median(iso)
                bytes_frame = RGBA8_BYTES_PER_PIXEL * n_pix
                literal_frame = {
                    "workload": WORKLOADS["frame"]["name"], "ms_per_frame_back_to_back": bb_ms,
                    "ms_per_frame_l2_flushed": iso_ms, "msamples_per_s_back_to_back": n_pix / (bb_ms * 1e-3) / 1e6,
                    "roofline": {"kernel": "csmain_frame_vec4", "bound": "fp32 (HBM figure for reference)", "unit": "GB/s",
                                 "achieved": bytes_frame / (iso_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"],
                                 "frac": bytes_frame / (iso_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                                 "algorithmic_bytes_per_launch": bytes_frame,
                                 "note": "8.3 MB written per launch = 1.3 us at the HBM roof; the back-to-back figure is dominated by "
                                         "launch latency (one ~5 us kernel per frame, nothing to overlap it with)"}}

    tot = torch.tensor([total_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(tot[0]), float(tot[1])

    samples_per_step = n_pix * spp_step                  # all ranks together
    value = samples_per_step * K / (total_ms * 1e-3) / 1e6
    e2e_value = samples_per_step * e2e_steps / (e2e_ms * 1e-3) / 1e6
    spp_rank = spp_step / world if args.shard != "tiles" else spp_step   # per-pixel samples one launch evaluates
    pix_rank = n_pix if args.shard != "tiles" else n_pix / world

    if rank == 0:
        # ---- roofline of the dominant kernel, from the live per-launch event durations -------------------
        k_ms = statistics.mean(kern_ms)
        if literal:
            alg = RGBA8_BYTES_PER_PIXEL * n_pix
            achieved = alg / (k_ms * 1e-3) / 1e9
            roof = {"kernel": "csmain_frame_vec4", "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": None,
                    "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']})",
                    "algorithmic_bytes_per_launch": alg, "launch_ms": k_ms}
        elif scene is not None:
            # algorithmic bytes per sample, COUNTED: 64 B per node visit, 16 B per sphere / 48 B per
            # triangle test, 2 x 40 B of path state per segment, 32 B accumulation read-modify-write per sample
            c = scene_counters
            prim_b = 16.0 if wl["scene"] == "spheres" else 48.0
            b_sample = (c["nodes_visited_per_sample"] * 64.0 + c["prims_tested_per_sample"] * prim_b +
                        c["segments_per_sample"] * 2.0 * STATE_BYTES_PER_PATH + ACCUM_BYTES_PER_PIXEL)
            n_launch = max(1.0, launches / K)                             # measured: kernels of this library per step
            alg = b_sample * pix_rank                                    # one sample of every pixel of the rank, all bounces
            wave_ms = k_ms / max(1.0, spp_rank)
            achieved = alg / (wave_ms * 1e-3) / 1e9
            # the part of those bytes that cannot be cache hits: path state streamed once per segment (evict-first, waves of
            # up to 32 Mi paths x 40 B), the per-sample radiance slot (16 B written, 16 B read back) and the accumulator
            stream_b = c["segments_per_sample"] * 2.0 * STATE_BYTES_PER_PATH + 32.0 + ACCUM_BYTES_PER_PIXEL / max(1.0, min(64.0, spp_rank))
            stream_gbs = stream_b * pix_rank / (wave_ms * 1e-3) / 1e9
            roof = {"kernel": "bounce_kernel", "bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"], "traffic": None,
                    "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']}); the scene ({c['scene_bytes'] / 1e6:.1f} MB) is "
                                   "L2-resident, so node/primitive fetch is served by L1/L2 (frac can exceed 1: it counts those cache hits "
                                   "as bytes, as SURVEY.md 8d prescribes) and the kernel is latency / divergence bound rather than HBM "
                                   "bound; hbm_stream_* is the share that must come from HBM (DESIGN.md 5)",
                    "algorithmic_bytes_per_sample": b_sample, "algorithmic_bytes_per_wave": alg,
                    "hbm_stream_bytes_per_sample": stream_b, "hbm_stream_gbs": stream_gbs, "hbm_stream_frac": stream_gbs / peaks["hbm_gbs"],
                    "launches_per_step": n_launch, "wave_ms": wave_ms, "launch_ms": k_ms / n_launch,
                    "note": "achieved = algorithmic bytes of ONE sample of every pixel / (render time per step / spp per step); "
                            "a wave carries several samples of every pixel (ptb_scene_set_wave_samples), launches_per_step counts "
                            "the bounce kernels and slice reductions actually launched",
                    "counters": c}
        else:
            fp32_peak = 148 * 128 * 2 * peaks["sm_max_mhz"] * 1e6 / 1e12     # TFLOP/s at max SM clock
            flops = FLOP_PER_SAMPLE * pix_rank * spp_rank
            achieved = flops / (k_ms * 1e-3) / 1e12
            roof = {"kernel": "csmain_accumulate", "bound": "fp32", "achieved": achieved, "peak": fp32_peak,
                    "unit": "TFLOP/s", "frac": achieved / fp32_peak, "traffic": None,
                    "peak_source": f"148 SM x 128 lanes x 2 x sm_max_mhz from MEASURED_PEAKS.json ({peaks['source']}); "
                                   "FP32 CUDA-core (FMA port) bound: no tensor or HBM roof applies (SURVEY.md 8d, DESIGN.md 4)",
                    "algorithmic_flops_per_launch": flops, "launch_ms": k_ms,
                    "hbm_bytes_per_launch": ACCUM_BYTES_PER_PIXEL * pix_rank,
                    "hbm_frac": ACCUM_BYTES_PER_PIXEL * pix_rank / (k_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr) and world == 1 and args.workload in ("c2", "c3", "c4"):
            try:
                key = roof["kernel"] if args.workload == "c2" else f"{roof['kernel']}:{args.workload}"
                roof["traffic"] = json.load(open(tr)).get(key)
            except Exception:
                pass

        # ---- CPU baseline on this box's host cores (bounded sample) ------------------------------------------
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            use_host_cores()
        if world == 1 and not args.no_cpu_baseline and scene is not None:
            ms, _, threads = cpu_scene_sample(wl, scene, 1, 2, 1)
            cpu = {"value": ms, "unit": "Msamples/s", "cores": threads, "kind": "port",
                   "sample": f"{W}x{H}, 1 of {spp} spp per step x 2 steps, scene oracle (own CPU BVH), OpenMP over rows, {cpu_model()}"}
        elif world == 1 and not args.no_cpu_baseline:
            spp_s = 1 if literal else max(1, min(spp, args.cpu_spp))
            ms, _, threads = cpu_sample(W, H, spp_s, 5, 1, jitter=not literal)
            cpu = {"value": ms, "unit": "Msamples/s", "cores": threads, "kind": "port",
                   "sample": f"{W}x{H}, {spp_s} of {spp} spp per step x 5 steps, OpenMP over rows, {cpu_model()}"}

        if world == 1:
            sharding = "none"
        elif args.shard == "tiles":
            sharding = "interleaved image rows per rank; merge = NCCL reduce of disjoint RGBA8 rows"
        elif args.shard == "p2p":
            sharding = ("disjoint sample ranges per rank; merge = one fused kernel per GPU over NVLink peer memory (P2P loads of "
                        "the partial accumulation buffers for the rank's pixel slab, rank-order sum, resolve, P2P store into rank 0's "
                        "image), double-buffered partials: the merge of frame k overlaps the rendering of frame k+1, drained inside "
                        "the timed total")
        else:
            sharding = "disjoint sample ranges per rank; merge = NCCL reduce(sum) of float4 accumulation buffers + resolve on rank 0"
        sharding += "" if world == 1 else (", every step" if merge_every_step else ", once, inside the timed total")
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": total_ms / K, "ms_per_frame": total_ms / K, "higher_is_better": True,
            "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["name"], "width": W, "height": H, "spp_per_step_total": spp_step,
                       "spp_per_step_per_rank": spp_rank, "seed": 0, "jitter": not literal, "sharding": sharding,
                       "step": ("ptb_render_resolve: accumulate + resolve in one launch" if fused else
                                "ptb_compute_frame" if literal else "ptb_render_samples + merge/resolve"),
                       "max_depth": wl.get("max_depth", 0), "scene": scene.name if scene is not None else "default (Test.hlsl literals)",
                       "l2": ("no flush: each step streams two 16 B/pixel partial buffers (> 126 MB L2) plus the peers' slabs"
                              if (pipelined and 2 * n_pix * 16 > 126 * 1024 * 1024) else
                              "flushed between timed steps (256 MB fill, " + ("inside the timed interval)" if pipelined else "untimed)")),
                       "timing": ("one CUDA-event interval over all K steps incl. the L2-flush fills and the drain of in-flight merges; max over ranks"
                                  if pipelined else "per-step CUDA events on the launching stream, summed; max over ranks")},
            "roofline": roof,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": 128, "d2h_bytes_per_step": n_pix * 4,
                    "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps,
                    "note": "C-ABI calls with host buffers (" + ("ptb_render_resolve = accumulate + resolve in one launch" if fused else
                                                                 "ptb_render_samples, ptb_resolve") + ", ptb_read_rgba8_async into "
                            "double-buffered pinned memory; merge over NCCL every step when N>1). The path's only input is "
                            "the parameter block (Test.hlsl reads nothing); output = the RGBA8 frame, every step"},
            "gpu_launches": int(launches),
            "kernel_mode": t.kernel_mode,
            "clocks": {"sm_mhz": clk["sm_mhz"], "sm_max_mhz": clk["sm_max_mhz"], "reasons": clk["reasons"],
                       "samples": clk["samples"], "power_w_max": clk["power_w_max"]},
            "merge_ms": merge_ms,
            "literal_frame": literal_frame,
        }
        print(json.dumps(line))
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel per step")
    ap.add_argument("--cpu-spp", type=int, default=16, help="spp per step of the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--shard", default=None, choices=["samples", "tiles", "p2p"],
                    help="multi-GPU partitioning: disjoint sample ranges merged by NCCL reduce (default), interleaved image "
                         "rows, or sample ranges merged by the fused peer-memory kernel (p2p)")
    args = ap.parse_args()
    if args.shard is None:
        # strong scaling merges every frame: the fused, pipelined peer-memory merge; otherwise sample ranges + NCCL reduce
        n = max(args.gpus, int(os.environ.get("WORLD_SIZE", "1")))
        args.shard = "samples"
        if n > 1 and args.workload == "c5":
            args.shard = "p2p"
        elif n > 1 and args.workload == "c4":
            args.shard = "tiles"                   # configs[3]: row-interleaved image tiles, bit-identical to one GPU
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
