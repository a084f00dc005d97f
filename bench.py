#!/usr/bin/env python
"""Headless benchmark of the path-tracing hot path (BASELINE.json: Mrays/s & ms/pass @1080p).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config sphere_grid] [--impl reference]

One "step" = one progressive 1-spp pass of the workload over the whole frame on each GPU. With N > 1 (torchrun, one
rank per GPU) the passes are sharded by sample index (rank r renders sampleIndex r, r+N, ...: SURVEY.md §8e) and the fp32
accumulation buffers are NCCL-reduced to rank 0 at the end of the timed region, so per-GPU work is fixed (weak scaling).
`--impl reference` times the CPU restatement of the reference's shaders (oracle/, all host threads) on the same config.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from heatray_b200 import capi, renderer, scenes  # noqa: E402

METRIC, UNIT = "path_tracing_throughput", "Mrays/s"


def load_traffic(config, kernel):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json), else None."""
    p = ROOT / "profiles" / "traffic.json"
    try:
        return json.loads(p.read_text())[config][kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks and throttle reasons while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.samples, self.stop_flag, self.proc, self.source = device, [], False, None, "nvidia-smi -lms 100"

    def run(self):
        if self._run_nvml():
            return
        try:
            proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                    stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.proc = proc
        for line in proc.stdout:
            if self.stop_flag:
                break
            parts = [x.strip() for x in line.split(",")]
            if len(parts) >= 6:
                self.samples.append(parts + [time.perf_counter()])
        try:
            proc.kill()
        except Exception:
            pass

    def _run_nvml(self):
        """NVML in-process (nvidia_ml_py): one sample every 10 ms, so even a 50 ms timed region is covered. False -> fall back to nvidia-smi."""
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(int(self.device))
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap))
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM); nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            return False
        self.source = "nvml, 10 ms"
        while not self.stop_flag:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
            except Exception:
                break
            self.samples.append([str(sm), str(mx)] + ["Active" if r & bit else "Not Active" for _, bit in bits] + [time.perf_counter()])
            time.sleep(0.01)
        return True

    def wait_ready(self, timeout=5.0):
        """nvidia-smi takes a few hundred ms to print its first line: wait for it so that a short timed region is not missed."""
        t = time.perf_counter()
        while not self.samples and time.perf_counter() - t < timeout:
            time.sleep(0.02)

    def finish(self, window=None):
        self.stop_flag = True
        try:
            self.proc.kill()
        except Exception:
            pass
        sm, mx, reasons = [], 0.0, set()
        for s in self.samples:
            try:
                sm.append(float(s[0])); mx = max(mx, float(s[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        out = {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}
        if window:
            out["samples_in_timed_region"] = sum(1 for x in self.samples if window[0] <= x[-1] <= window[1])
            out["source"] = self.source
            out["window"] = "samples from the start of the timed region to the end of the instrumented replay of the same passes (same kernels, same load)"
        return out


def build_spec(name, small=False):
    if name == "sphere_grid":
        return scenes.sphere_grid() if not small else scenes.sphere_grid(width=480, height=270, grid=4, u_slices=32, v_slices=16, env_size=(512, 256))
    if name == "material_grid":
        return scenes.material_grid() if not small else scenes.material_grid(width=480, height=270, grid=5, u_slices=32, v_slices=16)
    if name == "terrain":
        return scenes.terrain() if not small else scenes.terrain(width=960, height=540, n=256, blobs=20, env_size=(512, 256))
    if name == "cornell":
        return scenes.cornell_box()
    if name == "convergence":
        return scenes.sphere_grid(width=3840, height=2160, passes=1024) if not small else scenes.sphere_grid(width=960, height=540, passes=64, grid=4, u_slices=32, v_slices=16, env_size=(512, 256))
    raise SystemExit(f"unknown config {name}")


WORKLOAD_NAMES = {
    "sphere_grid": "configs[1]: procedural Sponza-scale 8x8 sphere grid, PBR roughness/metallic + env light + NEE",
    "material_grid": "configs[2]: 11x11 glass + clearcoat material grid, f/2.8 hexagonal bokeh, spot + directional",
    "terrain": "configs[3]: displaced terrain + flattened foliage",
    "cornell": "configs[0]: Cornell box, 1 point light",
    "convergence": "configs[4]: convergence run 3840x2160, 1024 spp sharded by sample index, NCCL reduce of the fp32 accumulation buffer",
}


def config_dict(spec, name, n_gpus, extra=None, shard="sample"):
    par = "single GPU" if n_gpus <= 1 else (f"sample-index sharding x{n_gpus}, one NCCL fp32 sum-reduce of the accumulation buffer" if shard == "sample"
                                            else f"screen-band sharding x{n_gpus} (interleaved 16-row bands), one NCCL fp32 sum-reduce of the disjoint bands")
    d = {"workload": WORKLOAD_NAMES[name], "scene": spec.name, "triangles": spec.triangle_count, "resolution": f"{spec.width}x{spec.height}",
         "max_ray_depth": spec.max_depth, "spp_per_step": 1, "qmc": "owen-scrambled sobol (16 sequences)",
         "parallelism": par,
         "l2_policy": "inputs larger than L2: every pass streams its ray / hit queues (%d MB at this resolution) through the 126 MB L2; the BVH + triangles are %s" % (
             spec.width * spec.height * 204 // 1000000, "L2-resident by design (< 126 MB)" if spec.triangle_count < 1000000 else "larger than L2 as well")}
    if extra:
        d.update(extra)
    return d


def algorithmic_bytes(st, closest=True):
    """B_ray = 32 + 80*N_nodes + 48*N_tris + 16 (closest) or + 4 (occlusion), summed over the rays of the launch."""
    if closest:
        rays = st.raysPrimary + st.raysExtension
        return 32 * rays + 80 * st.nodesVisited + 48 * st.trianglesTested + 16 * rays
    rays = st.raysOcclusion
    return 32 * rays + 80 * st.nodesVisitedAny + 48 * st.trianglesTestedAny + 4 * rays


def run_reference(args, rank, world):
    """CPU arm: the oracle (a restatement of the reference's shaders; OpenRL itself cannot run here), all host threads."""
    if rank != 0:
        return
    from oracle import loader
    spec = build_spec(args.config, args.small)
    b = loader.backend(fast=True)
    r = renderer.Renderer(b)
    r.load_scene(spec)
    cores = os.cpu_count() or 1
    bands = 16                      # one step = 1/16 of the frame rows (interleaved 8-row bands): a bounded sample of one pass
    rays = 0
    step_ms = []
    for i in range(args.warmup + args.steps):
        r.set_band(8, i % bands, bands)
        t0 = time.perf_counter()
        st = r.render_pass(sample_index=(i // bands) % spec.passes)
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            rays += st.rays
            step_ms.append(dt * 1e3)
    total_s = sum(step_ms) / 1e3
    value = rays / total_s / 1e6
    sample = f"{args.steps} steps, each 1/{bands} of the rows of one {spec.width}x{spec.height} pass (interleaved 8-row bands)"
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": statistics.mean(step_ms), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference", "config": config_dict(spec, args.config, 1, {"step": sample}),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "note": "CPU restatement of reference shaders + own BVH; OpenRL itself unavailable on this box"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline_sample(spec, budget_s=12.0):
    from oracle import loader
    b = loader.backend(fast=True)
    r = renderer.Renderer(b)
    r.load_scene(spec)
    bands, rays, t_total, done = 16, 0, 0.0, 0
    while t_total < budget_s or done < 2:          # whole bands of successive passes until the time budget is spent
        r.set_band(8, done % bands, bands)
        t0 = time.perf_counter()
        st = r.render_pass(sample_index=(done // bands) % spec.passes)
        t_total += time.perf_counter() - t0
        rays += st.rays
        done += 1
    b.close()
    return {"value": rays / t_total / 1e6, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
            "sample": f"{done} band-steps = {done / bands:.2f} passes of {spec.width}x{spec.height} (each step 1/{bands} of the rows, interleaved 8-row bands), {t_total:.1f} s",
            "ms_per_pass_extrapolated": t_total / done * bands * 1e3,
            "note": "CPU restatement of reference shaders + own BVH (oracle/, -O3 x86-64-v3 build); OpenRL itself unavailable on this box"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)    # 128 passes of config #2 = 0.4 s timed: several 100 ms clock samples fall inside it
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--config", default="sphere_grid", choices=list(WORKLOAD_NAMES))
    ap.add_argument("--small", action="store_true", help="reduced-size scene (debug only; not a valid bench number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--shard", default=None, choices=["sample", "bands"],
                    help="multi-GPU partition (SURVEY 8e): sample-index (default; weak scaling) or interleaved screen bands (default for terrain; strong scaling)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    lib = capi.load_library()
    spec = build_spec(args.config, args.small)
    shard = args.shard or ("bands" if args.config == "terrain" else "sample")
    b = capi.Backend(lib, "hrc", device=local_rank, collect_stats=False)
    r = renderer.Renderer(b)
    r.load_scene(spec)
    info = b.scene_info()
    K, W = args.steps, args.warmup
    if shard == "bands" and world > 1:
        # screen-band sharding: every rank renders every sample index for its interleaved 16-row bands (strong scaling)
        passes_total = max(spec.passes, K)
        r.set_band(16, rank, world)
    else:
        # sample-index sharding: rank r renders sample indices r, r + world, ... (wrapped into the QMC table length)
        passes_total = max(spec.passes, K * world)
    r.reset(passes=passes_total)
    if shard == "bands" and world > 1:
        r.set_band(16, rank, world)

    def sidx(step):
        return step % passes_total if (shard == "bands" and world > 1) else (rank + step * world) % passes_total

    acc = None
    if world > 1:   # the exchange step: one fp32 sum-reduce of the accumulation buffer, in place on the device pointer
        ptr = b.device_pixels()

        class _Arr:
            __cuda_array_interface__ = {"shape": (spec.height, spec.width, 4), "typestr": "<f4", "data": (ptr, False), "version": 2}
        acc = torch.as_tensor(_Arr(), device=f"cuda:{local_rank}")
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def reduce_timed():
        """NCCL reduce timed with CUDA events on torch's current stream (the pass stream has been synchronised before)."""
        if world <= 1:
            return 0.0
        ev0.record()
        dist.reduce(acc, dst=0, op=dist.ReduceOp.SUM)
        ev1.record()
        ev1.synchronize()
        return ev0.elapsed_time(ev1)

    # ---- warm-up -----------------------------------------------------------------------------
    for i in range(W):
        r.render_pass(want_stats=False, sample_index=sidx(i))
    reduce_timed()
    b.clear()

    # ---- traversal counters for the roofline numerator (same sample indices as the timed region; not timed) ----
    b.set_option(capi.OPT_TRAVERSAL_COUNTERS, 1)
    b.reset_stats()
    for i in range(K):
        r.enqueue_pass(sample_index=sidx(i))
    counters = b.get_stats()
    b.set_option(capi.OPT_TRAVERSAL_COUNTERS, 0)
    b.clear()
    rays_total = counters.rays

    # ---- timed region: exactly K passes, inputs resident in HBM, enqueued back to back on the context's stream and timed
    #      with CUDA events on that stream (no host synchronisation inside), then the reduce ------------------------------
    for i in range(2 * 4):                       # untimed: every lane captures its pass graph for this launch shape (counters off) before the clock starts
        r.enqueue_pass(sample_index=sidx(i))
    b.clear()
    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.wait_ready()
    barrier()
    sampler.samples.clear()                      # idle samples from before the region do not count
    t0 = time.perf_counter()
    b.timer_start()
    for i in range(K):
        r.enqueue_pass(sample_index=sidx(i))
    pass_ms = b.timer_stop()
    reduce_ms = reduce_timed()
    dev_ms = pass_ms + reduce_ms
    barrier()
    t1 = time.perf_counter()
    wall_ms = (t1 - t0) * 1e3

    # ---- instrumented replay of the same K passes: CUDA events between the kernels give each kernel's launch durations
    #      (their sum is within ~2 % of the clean region above; the shares and the roofline use these) ---------------------
    b.clear()
    b.set_option(capi.OPT_KERNEL_TIMING, 1)
    ms_closest = ms_any = ms_shade = ms_raygen = inst_ms = 0.0
    launches = 0
    for i in range(K):
        st = r.render_pass(want_stats=True, sample_index=sidx(i))
        ms_closest += st.msClosest; ms_any += st.msAny; ms_shade += st.msShade; ms_raygen += st.msRaygen; inst_ms += st.gpuMs
        launches += int(st.closestLaunches + st.anyLaunches + st.shadeLaunches) + 3     # + clear, raygen, finish
    b.set_option(capi.OPT_KERNEL_TIMING, 0)
    clocks = sampler.finish(window=(t0, t1))

    # ---- end-to-end region: the reference-facing call sequence with HOST buffers (pass + PixelPackBuffer readback) ----
    b.clear()
    barrier()
    t0 = time.perf_counter()
    checksum = 0.0
    for i in range(K):
        r.render_pass(want_stats=False, sample_index=sidx(i))
        px = b.map_pixels()                       # D2H of the RGBA32F accumulation buffer into pinned memory, every pass
        checksum += float(px[0, 0, 3])
        b.unmap_pixels()
    if world > 1:
        reduce_timed()
    barrier()
    e2e_serial_ms = (time.perf_counter() - t0) * 1e3

    # The same K steps the way hr::PassGenerator runs them when the caller keeps its next renderPass job queued (host_pass_generator.cpp:
    # look-ahead): pass i+1 is started into the backend's deferred buffer before pass i is read back, so the host copy of every pass
    # (still one full readback per step, each holding exactly i+1 samples) overlaps the rendering of the next one.
    lookahead = max(1, min(3, int(os.environ.get("HR_BENCH_LOOKAHEAD", "3"))))      # HRC_MAX_DEFERRED_PASSES = 3

    def pipelined_steps(n):
        total = 0.0
        r.frame.sampleIndex = sidx(0)
        b.enqueue_pass(r.frame)
        nxt = 1
        for i in range(n):
            while nxt < n and nxt - i <= lookahead:          # passes i+1 .. i+lookahead render while pass i is copied to the host
                r.frame.sampleIndex = sidx(nxt)
                b.enqueue_pass_deferred(r.frame)
                nxt += 1
            px = b.map_pixels()
            total += float(px[0, 0, 3])
            b.unmap_pixels()
            if i + 1 < n:
                b.commit_deferred()                          # the oldest pending pass: i+1
        return total

    pipelined_steps(max(8, W))                    # untimed warm-up: the deferred buffers (a ring of 4) are allocated on first use
    b.clear()
    barrier()
    t0 = time.perf_counter()
    alpha_sum = pipelined_steps(K)
    if world > 1:
        reduce_timed()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    assert alpha_sum == checksum, "the pipelined steps did not each deliver exactly their own pass count"

    # ---- gather over ranks (max time, summed rays) -------------------------------------------------
    if world > 1:
        t = torch.tensor([dev_ms, wall_ms, e2e_ms, e2e_serial_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms, e2e_ms, e2e_serial_ms = t.tolist()
        rt = torch.tensor([float(rays_total)], device="cuda", dtype=torch.float64)
        dist.all_reduce(rt, op=dist.ReduceOp.SUM)
        rays_all = rt.item()
    else:
        rays_all = float(rays_total)

    if rank == 0:
        step_ms = max(dev_ms, 0.0) / K
        value = rays_all / (dev_ms / 1e3) / 1e6
        peak, peak_src = load_peaks()
        bytes_closest = algorithmic_bytes(counters, True)
        bytes_any = algorithmic_bytes(counters, False)
        dominant_closest = ms_closest >= ms_any
        dom_bytes, dom_ms = (bytes_closest, ms_closest) if dominant_closest else (bytes_any, ms_any)
        n_launch = K * int(st.closestLaunches)
        achieved = dom_bytes / (dom_ms / 1e3) / 1e9
        rays_c = counters.raysPrimary + counters.raysExtension
        roof = {"bound": "hbm", "kernel": "k_trace<closest>" if dominant_closest else "k_trace<any>",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": load_traffic(args.config, "k_trace<closest>" if dominant_closest else "k_trace<any>"),
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": dom_bytes / n_launch, "avg_launch_ms": dom_ms / n_launch, "launches": n_launch,
                "per_ray": {"closest": {"rays": int(rays_c), "nodes": counters.nodesVisited / max(1, rays_c), "tris": counters.trianglesTested / max(1, rays_c),
                                        "bytes": bytes_closest / max(1, rays_c), "ms": ms_closest, "GBps": bytes_closest / (ms_closest / 1e3) / 1e9,
                                        "Mrays_per_s": rays_c / (ms_closest / 1e3) / 1e6},
                            "any": {"rays": int(counters.raysOcclusion), "nodes": counters.nodesVisitedAny / max(1, counters.raysOcclusion),
                                    "tris": counters.trianglesTestedAny / max(1, counters.raysOcclusion),
                                    "bytes": bytes_any / max(1, counters.raysOcclusion), "ms": ms_any, "GBps": bytes_any / (max(ms_any, 1e-9) / 1e3) / 1e9,
                                    "Mrays_per_s": counters.raysOcclusion / (max(ms_any, 1e-9) / 1e3) / 1e6}},
                "kernel_share_of_step": {"closest": ms_closest / inst_ms, "any": ms_any / inst_ms, "shade": ms_shade / inst_ms, "raygen": ms_raygen / inst_ms},
                "instrumented_ms_per_step": inst_ms / K,
                "timing": "value: K passes enqueued back to back, CUDA events on the launching stream; kernel durations: CUDA events between the kernels in a replay of the same K passes",
                "note": "scene BVH fits L2 for this config, so DRAM traffic is dominated by queue streams; see DESIGN.md"}
        frame_bytes = spec.width * spec.height * 16
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": step_ms,
                "higher_is_better": True, "scaling": "strong" if (shard == "bands" and world > 1) else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config_dict(spec, args.config, world, shard=shard, extra={"bvh_nodes": int(info.nodeCount), "bvh_build_s": float(info.buildSeconds),
                                                                 "scene_bytes": int(info.nodeBytes + info.triangleBytes + info.shadingBytes),
                                                                 "pass_pipelining": "value: the K passes are enqueued back to back (hrcEnqueuePass, the reference's offline mode) and rotate over %s sets of "
                                                                                    "queues / accumulation buffers, so the latency-bound tail of a pass overlaps the head of the next; "
                                                                                    "e2e: passes i+1 .. i+%d render into deferred buffers while pass i is read back (e2e.serial: one step at a time)" % (os.environ.get("HRC_TUNE_PASS_LANES", "4"), lookahead)}),
                "rays_per_step_per_gpu": rays_total / K, "wall_ms_per_step": wall_ms / K,
                "clocks": clocks, "roofline": roof,
                "e2e": {"value": rays_all / (e2e_ms / 1e3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms / K,
                        "h2d_bytes_per_step": C.sizeof(capi.FrameParams),
                        "d2h_bytes_per_step": frame_bytes if not (shard == "bands" and world > 1) else frame_bytes // world,   # band mode reads back only the rank's own rows
                        "api": "per step: hrcEnqueuePassDeferred (keeps %d passes ahead) + hrcReadPixels/hrcReleasePixels (this pass, pinned host buffer) + hrcCommitDeferred (the next pass) - "
                               "hr::PassGenerator::renderPass's look-ahead for a caller that keeps its next jobs queued; every step delivers its own full frame to the host" % lookahead,
                        "lookahead": lookahead,
                        "serial": {"value": rays_all / (e2e_serial_ms / 1e3) / 1e6, "ms_per_step": e2e_serial_ms / K,
                                   "api": "hrcRenderPass + hrcReadPixels/hrcReleasePixels, one step at a time (the reference viewer's loop, HeatrayRenderer.cpp:366-408)"}},
                "gpu_launches": launches + (0 if os.environ.get("HRC_TUNE_PASS_LANES") == "1" else 3)}       # + the lane joins (k_add_pixels) at the end of the timed region
        if world > 1:
            line["reduce_ms"] = reduce_ms
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_sample(spec)
        print(json.dumps(line), flush=True)
    b.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
