#!/usr/bin/env python3
"""bench.py -- Mrays/s of the per-pixel pass on BASELINE.json's configuration (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path through the C ABI
  python bench.py --impl reference [...]                          the reference shader on the host cores

A "step" is one 1920x1080 frame of the 128^3 terrain scene (BASELINE config 2: primary + shadow rays)
per GPU; the recorded poses are cycled.  A "ray" is one RayMarch or RayMarchShadow invocation as the
reference algorithm executes it (voxel.glsl:259,302), counted once per pose by the exact kernel's
counters (tests prove they equal the oracle's).  N > 1: one process per GPU (torchrun), every rank renders
its own pose's frame (weak scaling, grid replicated) and the frames are all-gathered over NCCL on a
side stream, overlapped with the next step's render; time is the max over ranks.
Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (scene, poses, width, height, flags)
    "c2": ("terrain128", ["terrain_inside", "terrain_low", "terrain_above"], 1920, 1080, 0),
    "c3": ("refraction128", ["refraction_close"], 1920, 1080, 0),
    "c4": ("terrain128_glass", ["glass_wall", "terrain_inside"], 3840, 2160, 0),
    "c1": ("glass_cube32", ["start"], 400, 400, 1),
}
WORKLOAD_DESC = {
    "c2": "_TERRAIN 128^3 chunk, 1920x1080, primary + shadow rays, 256^2 atlas (BASELINE configs[1])",
    "c3": "_REFRACTION 128^3 chunk, 1920x1080, <= 4 reflection/refraction bounces (BASELINE configs[2])",
    "c4": "_TERRAIN + glass walls 128^3 chunk, 3840x2160 multi-bounce (BASELINE configs[3])",
    "c1": "_GLASS_CUBE + _HIGH_PERFORMANCE + _COLOR_ONLY 32^3, 400x400 (BASELINE configs[0])",
}
I_STEP = 16          # algorithmic thread-instructions per reference DDA step


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--gather", default="nccl", choices=["nccl", "none"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------ the reference arm (CPU)
def strips(height, rows=12, every=4):
    """Bounded sample of a frame: every `every`-th strip of `rows` scanlines."""
    return [(y, min(rows, height - y)) for y in range(0, height, rows * every)]


def cpu_reference_frame(common, p, grid, atlas, sample, threads):
    """Times the reference shader (oracle/_ref) - or the oracle port when _ref was never built - on `sample`
    strips of the frame `p`.  Returns (seconds, kind)."""
    use_ref = common.ref_lib() is not None
    t0 = time.perf_counter()
    for (y0, h) in sample:
        q = p.copy()
        q.tile_y0, q.tile_h = y0, h
        if use_ref:
            common.ref_render(q, grid, atlas, nthreads=threads)
        else:
            common.oracle_render(q, grid, atlas, nthreads=threads)
    return time.perf_counter() - t0, ("reference" if use_ref else "port")


def sample_rays(common, p, grid, atlas, sample, threads):
    rays = steps = 0
    for (y0, h) in sample:
        q = p.copy()
        q.tile_y0, q.tile_h = y0, h
        c = common.oracle_render(q, grid, atlas, nthreads=threads).counters
        rays += c["primary_rays"] + c["secondary_rays"] + c["shadow_rays"]
        steps += c["march_steps"] + c["shadow_steps"]
    return rays, steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import common
    common.build_oracle()
    sc, poses, W, H, flags = WORKLOADS[args.workload]
    grid, n, atlas, ts = common.scene(sc)
    threads = os.cpu_count() or 1
    frames = [common.make_params(W, H, sc, pose, flags) for pose in poses]
    # Bounded sample: start from 1/4 of the scanlines and thin it until K + W steps fit in ~150 s on this box.
    every = 4
    dt, kind = cpu_reference_frame(common, frames[0], grid, atlas, strips(H, 12, every), threads)
    budget = 150.0 / max(1, args.steps + args.warmup)
    while dt > budget and every < H // 24:
        every *= 2
        dt /= 2.0
    sample = strips(H, 12, every)
    rays = [sample_rays(common, p, grid, atlas, sample, threads)[0] for p in frames]
    for i in range(args.warmup):
        cpu_reference_frame(common, frames[i % len(frames)], grid, atlas, sample, threads)
    total_t = 0.0
    total_r = 0
    for i in range(args.steps):
        dt, kind = cpu_reference_frame(common, frames[i % len(frames)], grid, atlas, sample, threads)
        total_t += dt
        total_r += rays[i % len(frames)]
    v = total_r / total_t / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s at 1080p terrain scene", "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / max(1, args.steps), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD_DESC[args.workload], "poses": poses, "frame": [W, H]},
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind,
                         "sample": "every %dth 12-row strip of each frame (1/%d of the scanlines), rays counted on the same strips" % (every, every)},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "samples": len(sm),
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import common
    import voxelraytracer_b200 as vrt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc, poses, W, H, flags = WORKLOADS[args.workload]
    grid, n, atlas, ts = common.scene(sc)
    ctx = vrt.Context(local)
    ctx.upload_grid(grid, n)
    ctx.upload_atlas(atlas, ts)
    if args.mode == "exact":
        flags |= vrt.FLAG_EXACT
    # each rank takes its own pose sequence (weak scaling: one frame per rank per step)
    frames = [common.make_params(W, H, sc, poses[(i + rank) % len(poses)], flags) for i in range(len(poses))]
    npix = W * H

    # reference-equivalent ray / step counts per pose (exact kernel's counters == oracle's, see tests)
    counts = []
    for p in frames:
        q = p.copy()
        q.flags = (p.flags | vrt.FLAG_EXACT | vrt.FLAG_COUNTERS)
        ctx.reset_counters()
        ctx.render(q)
        c = ctx.counters()
        q.flags = p.flags | vrt.FLAG_COUNTERS
        ctx.reset_counters()
        ctx.render(q)
        a = ctx.counters()
        counts.append({"rays": c["primary_rays"] + c["secondary_rays"] + c["shadow_rays"], "steps": c["march_steps"] + c["shadow_steps"],
                       "hits": c["hits"], "actual": a["actual_steps"]})
    launches0 = ctx.kernel_launches()

    stream = torch.cuda.Stream()
    comm = torch.cuda.Stream()
    fb = [torch.empty((H, W, 4), dtype=torch.uint8, device="cuda") for _ in range(2)]     # double-buffered frame
    gathered = torch.empty((world * H, W, 4), dtype=torch.uint8, device="cuda") if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")                        # > 126 MB L2
    do_gather = world > 1 and args.gather == "nccl"

    def one_step(i, ev0=None, ev1=None, do_flush=True):
        """flush (untimed) -> [ev0] render -> NCCL all-gather of the frames [ev1].  Returns the completion event."""
        p = frames[i % len(frames)]
        buf = fb[i % 2]
        with torch.cuda.stream(stream):
            if do_flush:
                flush.fill_(i & 0xFF)                # evict L2 between timed windows (window starts after it)
            if ev0 is not None:
                ev0.record(stream)
            ctx.render_to(p, buf.data_ptr(), stream.cuda_stream)
            if ev1 is not None and not do_gather:
                ev1.record(stream)
        done = torch.cuda.Event()
        if do_gather:
            rendered = torch.cuda.Event()
            rendered.record(stream)
            with torch.cuda.stream(comm):
                comm.wait_event(rendered)
                dist.all_gather_into_tensor(gathered, buf)
                if ev1 is not None:
                    ev1.record(comm)
                done.record(comm)
        else:
            done.record(stream)
        return done

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                              # samples cover warm-up + the timed region (frames are ~0.1 ms each)
    for i in range(args.warmup):
        stream.wait_event(one_step(i))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    # Timed region: K steps, each its own CUDA-event window [render (+ gather)], L2 flushed before each window.
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        stream.wait_event(one_step(i, ev[i][0], ev[i][1]))     # no overlap between steps in this pass
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    step_ms = [a.elapsed_time(b) for a, b in ev]
    t_steps = sum(step_ms) / 1e3

    # kernel-only windows (for the roofline): the same steps without the gather inside the window
    if do_gather:
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for i in range(args.steps):
            with torch.cuda.stream(stream):
                flush.fill_(i & 0xFF)
                kev[i][0].record(stream)
                ctx.render_to(frames[i % len(frames)], fb[i % 2].data_ptr(), stream.cuda_stream)
                kev[i][1].record(stream)
        torch.cuda.synchronize()
        t_kernel = sum(a.elapsed_time(b) for a, b in kev) / 1e3
    else:
        t_kernel = t_steps
    rays = sum(counts[i % len(frames)]["rays"] for i in range(args.steps))
    steps_ref = sum(counts[i % len(frames)]["steps"] for i in range(args.steps))
    steps_act = sum(counts[i % len(frames)]["actual"] for i in range(args.steps))
    hits = sum(counts[i % len(frames)]["hits"] for i in range(args.steps))

    # Pipelined pass (informational): the gather of frame i overlaps the render of frame i+1 on the comm stream;
    # no L2 fills here, so it is reported beside the headline, not as it.
    t_pipeline = None
    if do_gather:
        torch.cuda.synchronize(); dist.barrier()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pend = []
        p0.record(stream)
        for i in range(args.steps):
            if len(pend) >= 2:
                stream.wait_event(pend[-2])          # frame buffer i%2 was last gathered two steps ago
            pend.append(one_step(i, do_flush=False))
        stream.wait_event(pend[-1])
        p1.record(stream)
        torch.cuda.synchronize(); dist.barrier()
        t_pipeline = p0.elapsed_time(p1) / 1e3

    # e2e: the host-facing call (vrt_render_host): params in, frame back in pinned host memory, every step
    host, hp = ctx.host_alloc(npix * 4)
    for i in range(3):
        ctx.render_host(frames[i % len(frames)], hp, stream.cuda_stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        ctx.render_host(frames[i % len(frames)], hp, stream.cuda_stream)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    launches = ctx.kernel_launches() - launches0 - 3  # minus the e2e warm-ups

    # max over ranks
    t_step_total = t_steps
    if world > 1:
        v = torch.tensor([t_step_total, t_kernel, t_e2e, t_pipeline or 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        t_step_total, t_kernel_max, t_e2e, t_pipeline = [float(x) for x in v.cpu()]
        s = torch.tensor([rays, steps_ref, steps_act, hits], dtype=torch.float64, device="cuda")
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        rays_all, steps_all, act_all, hits_all = [float(x) for x in s.cpu()]
    else:
        t_kernel_max = t_kernel
        rays_all, steps_all, act_all, hits_all = float(rays), float(steps_ref), float(steps_act), float(hits)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        f_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        p_issue = sms * 4 * 32 * f_mhz * 1e6                          # thread-instructions / s at the sampled clock
        # dominant kernel = render_kernel; per launch (this rank): algorithmic instructions and bytes
        per_launch_s = t_kernel / args.steps
        instr_per_launch = I_STEP * steps_ref / args.steps
        bytes_per_launch = (steps_ref * 1 + hits * 4 + npix * 4 * args.steps) / args.steps
        ach_issue = instr_per_launch / per_launch_s
        line = {
            "metric": "Mrays/s at 1080p terrain scene" if args.workload == "c2" else "Mrays/s", "value": rays_all / t_step_total / 1e6,
            "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_step_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD_DESC[args.workload], "poses": poses, "frame": [W, H], "traversal": args.mode,
                       "sharding": "one frame per GPU per step (pose batch), grid replicated" + (", frames all-gathered over NCCL, overlapped" if do_gather else ""),
                       "l2": "flushed: 256 MiB fill before every timed step window (the fill itself is outside the window)",
                       "maxRayLength": 100.0},
            "frames_per_s": world * args.steps / t_step_total,
            "kernel_ms_per_frame": 1e3 * t_kernel_max / args.steps,
            "rays_per_frame": rays_all / (world * args.steps), "ref_steps_per_frame": steps_all / (world * args.steps),
            "actual_steps_per_frame": act_all / (world * args.steps),
            "gsteps_per_s": steps_all / t_step_total / 1e9,
            "roofline": {"bound": "issue", "achieved": ach_issue / 1e12, "peak": p_issue / 1e12, "unit": "Tinstr/s", "frac": ach_issue / p_issue,
                         "traffic": None, "peak_source": "148 SM x 4 schedulers x 32 lanes x sampled SM clock %.0f MHz" % f_mhz,
                         "algorithmic": "%d thread-instr per reference DDA step x reference steps (SURVEY.md 8(d))" % I_STEP},
            "roofline_hbm": {"bound": "hbm", "achieved": bytes_per_launch / per_launch_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                             "frac": bytes_per_launch / per_launch_s / 1e9 / hbm_peak, "traffic": None,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback"},
            "e2e": {"value": rays_all / t_e2e / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": C.sizeof(vrt.Params), "d2h_bytes_per_step": npix * 4,
                    "call": "vrt_render_host (params by value, RGBA8 frame copied to pinned host memory, synchronised)"},
            "gpu_launches": int(args.steps),          # one render_kernel launch per timed step (per rank)
            "gpu_launches_total": int(launches),      # every kernel this process's ctx launched after setup
            "pipelined": None if not t_pipeline else {"value": rays_all / t_pipeline / 1e6, "unit": "Mrays/s",
                                                       "note": "gather of frame i overlapped with render of frame i+1; no L2 fills"},
            "clocks": clocks,
            "wall_s": t_wall,
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            sample = strips(H)
            p = frames[0]
            r_s, _ = sample_rays(common, p, grid, atlas, sample, threads)
            dt, kind = cpu_reference_frame(common, p, grid, atlas, sample, threads)
            line["cpu_baseline"] = {"value": r_s / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": kind,
                                    "sample": "pose %s, every 4th 12-row strip (1/4 of the scanlines), %.1f s" % (poses[0], dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    ctx.close()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
