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
    if name == "sphere_grid_1m":
        return scenes.sphere_grid_1m() if not small else scenes.sphere_grid(width=480, height=270, grid=6, u_slices=32, v_slices=16, env_size=(512, 256), name="sphere_grid_1m")
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
    "sphere_grid_1m": "north_star target workload: ~1M-triangle scene (16x16 grid of the reference's 64x32 lat-long spheres), 1080p, 8 bounces, PBR + env light + NEE",
    "material_grid": "configs[2]: 11x11 glass + clearcoat material grid, f/2.8 hexagonal bokeh, spot + directional",
    "terrain": "configs[3]: displaced terrain + flattened foliage",
    "cornell": "configs[0]: Cornell box, 1 point light",
    "convergence": "configs[4]: convergence run 3840x2160, 1024 spp sharded by sample index, NCCL reduce of the fp32 accumulation buffer",
}
BAND_ROWS = int(os.environ.get("HR_BENCH_BAND_ROWS", "16"))
E2E_WARM = 8              # untimed caller-loop steps in front of every end-to-end region


def config_dict(spec, name, n_gpus, extra=None, shard="sample"):
    par = "single GPU" if n_gpus <= 1 else (f"sample-index sharding x{n_gpus}, one NCCL fp32 sum of the accumulation buffers" if shard == "sample"
                                            else f"screen-band sharding x{n_gpus} (interleaved {BAND_ROWS}-row bands), one NCCL sum onto rank 0 (the other ranks' rows are zeros)")
    d = {"workload": WORKLOAD_NAMES[name], "scene": spec.name, "triangles": spec.triangle_count, "resolution": f"{spec.width}x{spec.height}",
         "max_ray_depth": spec.max_depth, "spp_per_step": 1, "qmc": "owen-scrambled sobol (16 sequences)",
         "parallelism": par,
         "l2_policy": "inputs larger than L2: every pass streams its ray / hit queues (%d MB at this resolution) through the 126 MB L2; the BVH + triangles are %s" % (
             spec.width * spec.height * 204 // 1000000, "L2-resident by design (< 126 MB)" if spec.triangle_count < 1000000 else "not guaranteed to stay in L2 beside them")}
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
    """CPU arm: the oracle (a restatement of the reference's shaders, pinned to the shader text; OpenRL itself cannot run here), all host threads."""
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
                             "note": "CPU restatement of reference shaders (pinned bit for bit to the shader text, tests/test_shader_pinning.py) + own BVH; OpenRL itself unavailable on this box"},
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


class Rig:
    """One rank's backend + renderer + the torch view of its accumulation buffer and the exchange step of its sharding mode."""

    def __init__(self, lib, spec, shard, rank, world, local_rank, steps):
        import torch
        self.torch, self.spec, self.shard, self.rank, self.world, self.local_rank = torch, spec, shard, rank, world, local_rank
        self.bands = shard == "bands" and world > 1
        self.b = capi.Backend(lib, "hrc", device=local_rank, collect_stats=False)
        self.r = renderer.Renderer(self.b)
        self.r.load_scene(spec)
        self.info = self.b.scene_info()
        # sample-index sharding: rank r renders sample indices r, r + world, ... (wrapped into the QMC table length);
        # screen-band sharding: every rank renders every sample index for its interleaved bands (strong scaling)
        # the e2e region renders E2E_WARM + steps consecutive passes of one progressive render, and the look-ahead must still find passes to predict at its end
        self.passes_total = max(spec.passes, (steps + E2E_WARM + 4) * (1 if self.bands else world))
        if self.bands:
            self.r.set_band(BAND_ROWS, rank, world)
        self.r.reset(passes=self.passes_total)
        if self.bands:
            self.r.set_band(BAND_ROWS, rank, world)
        self.acc = None
        if world > 1:
            ptr = self.b.device_pixels()

            class _Arr:
                __cuda_array_interface__ = {"shape": (spec.height, spec.width, 4), "typestr": "<f4", "data": (ptr, False), "version": 2}
            self.acc = torch.as_tensor(_Arr(), device=f"cuda:{local_rank}")
            self.red = torch.empty_like(self.acc)                 # the exchange step of the e2e region works on a copy: the local sum keeps growing
            self.host = torch.empty(self.acc.shape, dtype=self.acc.dtype, pin_memory=True) if rank == 0 else None
        self.peer = None
        if world > 1 and self.collective() == "peer":
            from heatray_b200 import sharding
            self.peer = sharding.PeerExchange(lib, local_rank, rank, world, spec.height, spec.width)
        self.ev0, self.ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def sidx(self, step):
        return step % self.passes_total if self.bands else (self.rank + step * self.world) % self.passes_total

    def exchange(self, buf=None):
        """The one collective of the path, timed with CUDA events on torch's current stream (the pass streams were synchronised before).
        One fp32 sum of the accumulation buffers onto rank 0 (see exchange_async for the alternatives that were measured)."""
        if self.world <= 1:
            return 0.0
        import torch.distributed as dist
        from heatray_b200 import sharding
        buf = self.acc if buf is None else buf
        self.ev0.record()
        self.exchange_async(buf)
        self.ev1.record()
        self.ev1.synchronize()
        return self.ev0.elapsed_time(self.ev1)

    def collective(self):
        return os.environ.get("HR_BENCH_COLLECTIVE", "reduce")

    def combined(self, buf=None):
        """Rank 0: where the exchanged frame is (the peer exchange has its own result buffer, the NCCL collectives work in place)."""
        return self.peer.result() if self.peer else (self.acc if buf is None else buf)

    def exchange_async(self, buf):
        """Enqueues the exchange on torch's current stream (no host wait). HR_BENCH_COLLECTIVE: reduce (default: NCCL fp32 sum onto rank 0, in-switch
        reduction on NVSwitch; for bands the other ranks' rows are zeros, so the sum is exact), all_reduce, gather (band mode: each rank sends only its own
        rows), or peer: ONE own kernel per rank over NVLink peer memory (hrcExchangeSum / hrcExchangeBands, csrc/peer_exchange.cu). Measured back to back at
        8 ranks (profiles/r02_notes.md): 33 MB reduce 72 us, peer sum 130 us, peer bands 71 us; 133 MB reduce 221, peer bands 195 us."""
        import torch.distributed as dist
        from heatray_b200 import sharding
        kind = self.collective()
        if self.peer:
            stream = self.torch.cuda.current_stream().cuda_stream
            if self.bands:
                self.peer.bands(buf.data_ptr(), BAND_ROWS, stream)
            else:
                self.peer.sum(buf.data_ptr(), stream)
            return
        if self.bands and kind == "gather":
            sharding.gather_bands(buf, BAND_ROWS, self.rank, self.world, dst=0)
        elif kind == "all_reduce":
            dist.all_reduce(buf, op=dist.ReduceOp.SUM)
        else:
            dist.reduce(buf, dst=0, op=dist.ReduceOp.SUM)

    def close(self):
        if self.peer:
            self.peer.status()
            self.peer.close()
        self.b.close()


def measure_value(rig, K, barrier, sampler=None):
    """Exactly K passes, inputs resident in HBM, enqueued back to back (hrcEnqueuePass) and timed with CUDA events on the launching stream, then the exchange step."""
    b, r = rig.b, rig.r
    for i in range(2 * 4):                       # untimed: every lane captures its pass graph for this launch shape before the clock starts
        r.enqueue_pass(sample_index=rig.sidx(i))
    b.clear()
    b.reset_stats()                              # ray totals read after the region count its K passes only
    barrier()
    if sampler:
        sampler.samples.clear()                  # idle samples from before the region do not count
    t0 = time.perf_counter()
    b.timer_start()
    for i in range(K):
        r.enqueue_pass(sample_index=rig.sidx(i))
    pass_ms = b.timer_stop()
    ex_ms = rig.exchange()
    barrier()
    t1 = time.perf_counter()
    return pass_ms, ex_ms, t0, t1


def exchange_check(rig, barrier, passes_per_rank=4):
    """Multi-GPU correctness on the hardware: the exchanged frame of `passes_per_rank` passes per rank against rank 0's own single-GPU render of the
    same sample indices (sample mode: alpha exact, rgb up to fp32 summation order; band mode: bit-identical)."""
    import torch
    b, r, world = rig.b, rig.r, rig.world
    b.clear()
    for i in range(passes_per_rank):
        r.render_pass(want_stats=False, sample_index=rig.sidx(i))
    b.synchronize()
    barrier()
    rig.exchange()
    out = None
    if rig.rank == 0:
        multi = rig.combined().clone()
        b.clear()
        if rig.bands:
            r.set_band(0, 0, 1)
            indices = [rig.sidx(i) for i in range(passes_per_rank)]
        else:
            indices = [(q + i * world) % rig.passes_total for i in range(passes_per_rank) for q in range(world)]
        for s in indices:
            r.render_pass(want_stats=False, sample_index=s)
        b.synchronize()
        single = rig.acc
        alpha_exact = bool(torch.equal(multi[..., 3], single[..., 3])) and bool((single[..., 3] == len(indices)).all())
        denom = torch.clamp(single[..., :3].abs(), min=1e-3)
        max_rel = float(((multi[..., :3] - single[..., :3]).abs() / denom).max())
        out = {"alpha_exact": alpha_exact, "max_rel": max_rel, "bit_identical": bool(torch.equal(multi, single)),
               "passes": len(indices), "what": "exchanged N-rank frame vs rank 0's single-GPU render of the same sample indices"}
        if rig.bands:
            r.set_band(BAND_ROWS, rig.rank, world)
    b.clear()
    barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=128)    # 128 passes of config #2 = 0.4 s timed: several clock samples fall inside it
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--config", default="sphere_grid", choices=list(WORKLOAD_NAMES))
    ap.add_argument("--small", action="store_true", help="reduced-size scene (debug only; not a valid bench number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-strong", action="store_true", help="multi-GPU runs: skip the extra strong-scaling measurement of configs[3] (terrain, screen bands)")
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
        # the collective's kernel goes on a high-priority stream: with passes in flight on every lane a normal-priority NCCL kernel waits behind them for SMs
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=os.environ.get("HR_BENCH_NCCL_PRIORITY", "1") != "0")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), pg_options=opts)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    lib = capi.load_library()
    spec = build_spec(args.config, args.small)
    shard = args.shard or ("bands" if args.config == "terrain" else "sample")
    K, W = args.steps, args.warmup
    rig = Rig(lib, spec, shard, rank, world, local_rank, K)
    b, r, info, sidx = rig.b, rig.r, rig.info, rig.sidx
    bands = rig.bands

    # ---- warm-up -----------------------------------------------------------------------------
    for i in range(W):
        r.render_pass(want_stats=False, sample_index=sidx(i))
    rig.exchange()
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

    # ---- timed region ------------------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.wait_ready()
    pass_ms, reduce_ms, t0, t1 = measure_value(rig, K, barrier, sampler)
    dev_ms = pass_ms + reduce_ms
    wall_ms = (t1 - t0) * 1e3

    # ---- instrumented replay of the same K passes: CUDA events between the kernels give each kernel's launch durations
    #      (the shares and the roofline use these) ----------------------------------------------------------------------
    b.clear()
    b.set_option(capi.OPT_KERNEL_TIMING, 1)
    ms_closest = ms_any = ms_shade = ms_raygen = ms_primary = inst_ms = 0.0
    launches = 0
    for i in range(K):
        st = r.render_pass(want_stats=True, sample_index=sidx(i))
        ms_closest += st.msClosest; ms_any += st.msAny; ms_shade += st.msShade; ms_raygen += st.msRaygen; ms_primary += st.msClosestPrimary; inst_ms += st.gpuMs
        launches += int(st.closestLaunches + st.anyLaunches + st.shadeLaunches) + 2     # + raygen, finish
        if int(os.environ.get("HRC_TUNE_TAIL_RAYS", "0")) > 0:                       # the timed passes also launch k_tail_decide + k_path per iteration boundary
            launches += 2 * (int(st.closestLaunches) - 1)
    b.set_option(capi.OPT_KERNEL_TIMING, 0)
    clocks = sampler.finish(window=(t0, t1))

    # ---- end-to-end regions: the reference-facing call sequence with HOST buffers ---------------------------------------
    # The caller is the reference viewer's loop (HeatrayRenderer.cpp:366-408): hrcRenderPass, then map the pixel pack buffer, one pass at a
    # time. N > 1: every step ends with the exchange step on a copy of the accumulation buffers and rank 0 alone copies the combined frame
    # to its host (the other ranks have nothing to show).
    side = torch.cuda.Stream(priority=-1) if world > 1 else None
    snap = torch.cuda.Event() if world > 1 else None

    def caller_loop(first, n, overlapped=False):
        total = 0.0
        pending = None
        for i in range(first, first + n):
            r.render_pass(want_stats=False, sample_index=sidx(i))
            if world > 1 and overlapped:
                # the exchange of pass i (snapshot of the local sums -> collective -> rank 0's host copy) runs on a side stream while pass i+1 renders:
                # the combined frame reaches the host one pass late, every step still delivers one
                if pending is not None:
                    pending.synchronize()
                    if rank == 0:
                        total += float(rig.host[0, 0, 3])
                with torch.cuda.stream(side):
                    rig.red.copy_(rig.acc)
                    snap.record()
                    rig.exchange_async(rig.red)
                    if rank == 0:
                        rig.host.copy_(rig.combined(rig.red), non_blocking=True)
                    pending = torch.cuda.Event(); pending.record()
                snap.synchronize()                # the next pass may write the accumulation buffer once the snapshot is taken
            elif world > 1:
                if rig.peer:
                    rig.exchange(rig.acc)
                else:
                    rig.red.copy_(rig.acc)
                    rig.exchange(rig.red)
                if rank == 0:
                    rig.host.copy_(rig.combined(rig.red), non_blocking=True)
                    torch.cuda.synchronize()
                    total += float(rig.host[0, 0, 3])
            else:
                px = b.map_pixels()               # D2H of the RGBA32F accumulation buffer into pinned memory, every pass
                total += float(px[0, 0, 3])
                b.unmap_pixels()
        if pending is not None:
            pending.synchronize()
            if rank == 0:
                total += float(rig.host[0, 0, 3])
        return total

    def timed_caller_loop(spec_depth, overlapped=False):
        b.set_option(capi.OPT_SPECULATIVE_STRIDE, 1 if bands else world)      # a sample-sharded rank renders every world-th sample index
        b.set_option(capi.OPT_SPECULATIVE_PASSES, spec_depth)
        b.clear()
        # untimed: side buffers and their pass graphs exist before the clock starts. The timed steps then CONTINUE this progressive render (no clear in
        # between): a window of K steps of a long render, in which the passes the backend has running ahead at the window's start were paid for before it and
        # the ones running ahead at its end are paid for inside it — exactly K passes of GPU work. (Clearing here would throw the look-ahead away and
        # charge the window K + 3 passes: +15 % at K = 20.)
        caller_loop(0, E2E_WARM, overlapped)
        barrier()
        t = time.perf_counter()
        chk = caller_loop(E2E_WARM, K, overlapped)
        barrier()
        ms = (time.perf_counter() - t) * 1e3
        b.set_option(capi.OPT_SPECULATIVE_PASSES, 0)
        return ms, chk

    e2e_plain_ms, chk_plain = timed_caller_loop(0)
    e2e_spec_ms, chk_spec = timed_caller_loop(3)
    assert chk_plain == chk_spec, "speculated passes leaked into a frame the caller read"
    # One GPU: the headline is the caller loop with the backend's speculative look-ahead. Several GPUs: every step ends in a collective that all ranks
    # must reach together (passes speculated ahead on every rank delay it badly), so the headline overlaps the exchange of pass i with pass i+1 instead.
    e2e_overlap_ms = e2e_spec_overlap_ms = 0.0
    if world > 1:
        e2e_overlap_ms, chk_ov = timed_caller_loop(0, overlapped=True)
        e2e_spec_overlap_ms, chk_sov = timed_caller_loop(3, overlapped=True)
        assert chk_ov == chk_plain and chk_sov == chk_plain
    e2e_ms = e2e_spec_ms if world == 1 else e2e_spec_overlap_ms

    # The same K steps with the caller queueing ahead explicitly (hr::PassGenerator's look-ahead when renderPass jobs are queued behind the
    # running one): passes i+1 .. i+3 render into deferred buffers while pass i is read back.
    e2e_look_ms = None
    if world == 1:
        lookahead = 3

        def pipelined_steps(n):
            total = 0.0
            r.frame.sampleIndex = sidx(0)
            b.enqueue_pass(r.frame)
            nxt = 1
            for i in range(n):
                while nxt < n and nxt - i <= lookahead:
                    r.frame.sampleIndex = sidx(nxt)
                    b.enqueue_pass_deferred(r.frame)
                    nxt += 1
                px = b.map_pixels()
                total += float(px[0, 0, 3])
                b.unmap_pixels()
                if i + 1 < n:
                    b.commit_deferred()
            return total

        b.clear()
        pipelined_steps(max(8, W))
        b.clear()
        barrier()
        t = time.perf_counter()
        chk_look = pipelined_steps(K)
        barrier()
        e2e_look_ms = (time.perf_counter() - t) * 1e3
        assert chk_look == K * (K + 1) / 2, "a frame the caller read did not hold exactly its own sample count"      # this region starts from a clear: frame i holds i + 1 samples

    # ---- multi-GPU: correctness of the exchange on the hardware, and the strong-scaling case BASELINE.json names (configs[3] by bands) ----
    reduce_check = exchange_check(rig, barrier) if world > 1 else None
    exchange_name = (rig.collective() + (" (one kernel per rank over NVLink peer memory: hrcExchangeSum / hrcExchangeBands)" if rig.peer else " (NCCL)")) if world > 1 else None
    strong = None
    if world > 1 and not args.no_strong and args.config != "terrain" and not args.small:
        rig.close()
        rig = None
        strong = strong_terrain(lib, rank, world, local_rank, barrier)

    # ---- gather over ranks (max time, summed rays) -------------------------------------------------
    if world > 1:
        t = torch.tensor([dev_ms, wall_ms, e2e_ms, e2e_plain_ms, reduce_ms, e2e_spec_ms, e2e_overlap_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms, e2e_ms, e2e_plain_ms, reduce_ms, e2e_spec_ms, e2e_overlap_ms = t.tolist()
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
        dom_name = "k_trace<closest>" if dominant_closest else "k_trace<any>"
        n_launch = K * int(st.closestLaunches)
        achieved = dom_bytes / (dom_ms / 1e3) / 1e9
        rays_c = counters.raysPrimary + counters.raysExtension
        # secondary (extension) rays only: iterations >= 1 of the closest-hit kernel — the "incoherent secondary-ray throughput" of north_star
        sec_rays = counters.raysExtension
        sec_nodes = counters.nodesVisited - counters.nodesVisitedPrimary
        sec_tris = counters.trianglesTested - counters.trianglesTestedPrimary
        sec_ms = ms_closest - ms_primary
        sec_bytes = 48 * sec_rays + 80 * sec_nodes + 48 * sec_tris
        traffic = load_traffic(args.config, dom_name)
        avg_launch_ms = dom_ms / n_launch
        roof = {"bound": "hbm", "kernel": dom_name,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic,
                "peak_source": peak_src,
                "algorithmic_GBps": achieved,
                "dram_GBps_measured": (traffic / (avg_launch_ms / 1e3) / 1e9) if traffic else None,
                "dram_frac_measured": (traffic / (avg_launch_ms / 1e3) / 1e9 / peak) if traffic else None,
                "limiter": "issue slots, not DRAM: `frac` is the algorithmic node + triangle bytes per second against the HBM copy peak as the metric defines it; the BVH of this config is served "
                           "from L1 / L2 (ncu: issue-active 62-78 %, 14-21 of 32 lanes per instruction, DRAM 2-6 % of peak; profiles/*_ncu_trace_full_summary.txt), "
                           "dram_frac_measured is the kernel's real share of HBM bandwidth",
                "algorithmic_bytes_per_launch": dom_bytes / n_launch, "avg_launch_ms": avg_launch_ms, "launches": n_launch,
                "per_ray": {"closest": {"rays": int(rays_c), "nodes": counters.nodesVisited / max(1, rays_c), "tris": counters.trianglesTested / max(1, rays_c),
                                        "bytes": bytes_closest / max(1, rays_c), "ms": ms_closest, "GBps": bytes_closest / (ms_closest / 1e3) / 1e9,
                                        "Mrays_per_s": rays_c / (ms_closest / 1e3) / 1e6},
                            "secondary": {"rays": int(sec_rays), "nodes": sec_nodes / max(1, sec_rays), "tris": sec_tris / max(1, sec_rays),
                                          "bytes": sec_bytes / max(1, sec_rays), "ms": sec_ms, "GBps": sec_bytes / (max(sec_ms, 1e-9) / 1e3) / 1e9,
                                          "Mrays_per_s": sec_rays / (max(sec_ms, 1e-9) / 1e3) / 1e6,
                                          "what": "extension rays only: wavefront iterations >= 1 of k_trace<closest> (per GPU), launch durations from the instrumented replay"},
                            "any": {"rays": int(counters.raysOcclusion), "nodes": counters.nodesVisitedAny / max(1, counters.raysOcclusion),
                                    "tris": counters.trianglesTestedAny / max(1, counters.raysOcclusion),
                                    "bytes": bytes_any / max(1, counters.raysOcclusion), "ms": ms_any, "GBps": bytes_any / (max(ms_any, 1e-9) / 1e3) / 1e9,
                                    "Mrays_per_s": counters.raysOcclusion / (max(ms_any, 1e-9) / 1e3) / 1e6}},
                "kernel_share_of_step": {"closest": ms_closest / inst_ms, "any": ms_any / inst_ms, "shade": ms_shade / inst_ms, "raygen": ms_raygen / inst_ms},
                "instrumented_ms_per_step": inst_ms / K,
                "timing": "value: K passes enqueued back to back, CUDA events on the launching stream; kernel durations: CUDA events between the kernels in a replay of the same K passes",
                "note": "scene BVH fits L2 for this config, so DRAM traffic is dominated by queue streams; see DESIGN.md" if spec.triangle_count < 1000000 else "see DESIGN.md"}
        frame_bytes = spec.width * spec.height * 16
        e2e = {"value": rays_all / (e2e_ms / 1e3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms / K,
               "h2d_bytes_per_step": C.sizeof(capi.FrameParams),
               "d2h_bytes_per_step": frame_bytes,
               "api": ("per step: hrcRenderPass, then hrcReadPixels / hrcReleasePixels (pinned host buffer), one pass at a time - the reference viewer's loop (HeatrayRenderer.cpp:366-408) - "
                       "with HRC_OPT_SPECULATIVE_PASSES = 3: the backend keeps the next three passes of the progressive render running while the caller reads this one back; "
                       "every step delivers its own full frame (exactly i + 1 samples) to the host") if world == 1 else
                      ("per step and rank: hrcRenderPass with HRC_OPT_SPECULATIVE_PASSES = 3, then on a high-priority side stream a device snapshot of the accumulation "
                       "buffer, the exchange step on the snapshot and rank 0's copy of the combined frame to its pinned host buffer, overlapped with the next pass "
                       "(the combined frame reaches the host one pass late; the other ranks read nothing back); e2e.serial is the same loop with neither"),
               "speculative_passes": 3,
               "window": f"steps {E2E_WARM} .. {E2E_WARM} + K - 1 of ONE progressive render whose first {E2E_WARM} steps ran untimed (steady state: three passes run ahead at its start and at its end, the clock stops after the device has drained: exactly K passes of GPU work, K frames delivered)",
               "serial": {"value": rays_all / (e2e_plain_ms / 1e3) / 1e6, "ms_per_step": e2e_plain_ms / K,
                          "api": "the same caller loop with HRC_OPT_SPECULATIVE_PASSES = 0: one pass in flight, readback not overlapped"},
               "speculative3": {"value": rays_all / (e2e_spec_ms / 1e3) / 1e6, "ms_per_step": e2e_spec_ms / K,
                                "api": "the same caller loop with HRC_OPT_SPECULATIVE_PASSES = 3"}}
        if world > 1:
            e2e["overlapped_exchange"] = {"value": rays_all / (e2e_overlap_ms / 1e3) / 1e6, "ms_per_step": e2e_overlap_ms / K,
                                          "api": "the exchange overlapped with the next pass, HRC_OPT_SPECULATIVE_PASSES = 0"}
            e2e["speculative3"]["api"] += ", exchange not overlapped"
        if e2e_look_ms:
            e2e["lookahead3"] = {"value": rays_all / (e2e_look_ms / 1e3) / 1e6, "ms_per_step": e2e_look_ms / K,
                                 "api": "caller queues ahead itself: hrcEnqueuePassDeferred keeps 3 passes ahead, hrcReadPixels, hrcCommitDeferred (hr::PassGenerator with renderPass jobs queued)"}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": step_ms,
                "higher_is_better": True, "scaling": "strong" if bands else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config_dict(spec, args.config, world, shard=shard, extra={"bvh_nodes": int(info.nodeCount), "bvh_build_s": float(info.buildSeconds),
                                                                 "scene_bytes": int(info.nodeBytes + info.triangleBytes + info.shadingBytes),
                                                                 "pass_pipelining": "value: the K passes are enqueued back to back (hrcEnqueuePass, the reference's offline mode) and rotate over %s sets of "
                                                                                    "queues / accumulation buffers, so the latency-bound tail of a pass overlaps the head of the next; each lane replays its pass as one captured CUDA graph" % os.environ.get("HRC_TUNE_PASS_LANES", "8")}),
                "rays_per_step_per_gpu": rays_total / K, "wall_ms_per_step": wall_ms / K, "timed_region_ms": dev_ms,
                "clocks": clocks, "roofline": roof, "e2e": e2e,
                "gpu_launches": launches + 3}       # + the lane joins (k_add_pixels) at the end of the timed region
        if world > 1:
            line["reduce_ms"] = reduce_ms
            line["exchange"] = exchange_name
            line["reduce_check"] = reduce_check
            if strong:
                line["strong"] = strong
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_sample(spec)
        print(json.dumps(line), flush=True)
    if rig:
        rig.close()
    if world > 1:
        dist.destroy_process_group()


def strong_terrain(lib, rank, world, local_rank, barrier, steps=32):
    """configs[3] (terrain, 10 M triangles, 4K) by interleaved screen bands on all ranks: strong scaling. Also times the same passes on rank 0
    alone over the full frame (the other ranks wait), so that the line carries its own single-GPU figure."""
    import torch
    import torch.distributed as dist
    spec = build_spec("terrain")
    rig = Rig(lib, spec, "bands", rank, world, local_rank, steps)
    b, r = rig.b, rig.r
    for i in range(3):
        r.render_pass(want_stats=False, sample_index=rig.sidx(i))
    rig.exchange()                                # untimed: NCCL sets its gather connections up on first use
    b.clear()
    pass_ms, ex_ms, _, _ = measure_value(rig, steps, barrier)
    rays = float(b.get_stats().rays)
    check = exchange_check(rig, barrier, passes_per_rank=2)
    t = torch.tensor([pass_ms + ex_ms, ex_ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    rt = torch.tensor([rays], device="cuda", dtype=torch.float64)
    dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    out = None
    n1_ms = n1_rays = 0.0
    if rank == 0:                                  # the single-GPU figure: full frame, same passes, the other ranks idle
        r.set_band(0, 0, 1)
        b.clear(); b.reset_stats()
        for i in range(8):
            r.enqueue_pass(sample_index=i)
        b.clear(); b.reset_stats()
        b.timer_start()
        for i in range(steps):
            r.enqueue_pass(sample_index=i)
        n1_ms = b.timer_stop()
        n1_rays = float(b.get_stats().rays)
    barrier()
    if rank == 0:
        value = rt.item() / (t[0].item() / 1e3) / 1e6
        n1_value = n1_rays / (n1_ms / 1e3) / 1e6
        out = {"config": f"configs[3] terrain ({spec.triangle_count} triangles, {spec.width}x{spec.height}), interleaved {BAND_ROWS}-row screen bands, exchange = {rig.collective()}",
               "value": value, "unit": UNIT, "ms_per_step": t[0].item() / steps, "steps": steps, "exchange_ms": t[1].item(),
               "n1_value": n1_value, "n1_ms_per_step": n1_ms / steps, "speedup": value / n1_value, "efficiency": value / n1_value / world,
               "reduce_check": check}
    rig.close()
    return out


if __name__ == "__main__":
    main()
