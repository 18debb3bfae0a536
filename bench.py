#!/usr/bin/env python3
"""bench.py -- Mrays/s of the per-pixel pass on BASELINE.json's configuration (see DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path through the C ABI
  python bench.py --impl reference [...]                          the reference shader on the host cores

A "step" is one 1920x1080 frame of the 128^3 terrain scene (BASELINE config 2: primary + shadow rays)
per GPU; the recorded poses are cycled.  A "ray" is one RayMarch or RayMarchShadow invocation as the
reference algorithm executes it (voxel.glsl:259,302), counted once per pose by the exact kernel's
counters (tests prove they equal the oracle's).  N > 1: one process per GPU (torchrun), grid replicated.
  --scaling weak (default): every rank renders its own pose's frame - independent units, no data-path collective in the timed
      steps; a second pass measures the fused NVLink gather of ALL frames into rank 0 and reports it as `gathered`.
  --scaling strong: ONE frame per step, 16-row bands dealt round-robin over the ranks, every rank's render kernel storing its
      bands straight into rank 0's frame over NVLink (fused render + gather, CUDA IPC mapping).
Time is the max over ranks.  Prints ONE JSON line on rank 0.
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
    # C5: 64 poses on a circle of radius 600 around the 1024^3 chunk, looking inward; maxRayLength 2048
    "c5": ("synthetic1024", ["orbit%d" % k for k in range(64)], 7680, 4320, 0),
    "c5s": ("synthetic512", ["orbit%d" % k for k in range(64)], 3840, 2160, 0),   # half-size rehearsal of C5
}
MAX_LEN = {"c5": 2048.0, "c5s": 1024.0}


def pose_of(workload, name):
    """Named pose -> what common.make_params takes (orbit poses are generated)."""
    if name.startswith("orbit"):
        import common
        scale = 0.5 if workload == "c5s" else 1.0
        return common.orbit_pose(int(name[5:]), 64, 600.0 * scale, 250.0 * scale)
    return name


def mk(common, workload, W, H, sc, pose, flags):
    return common.make_params(W, H, sc, pose_of(workload, pose), flags, max_len=MAX_LEN.get(workload))

WORKLOAD_DESC = {
    "c2": "_TERRAIN 128^3 chunk, 1920x1080, primary + shadow rays, 256^2 atlas (BASELINE configs[1])",
    "c3": "_REFRACTION 128^3 chunk, 1920x1080, <= 4 reflection/refraction bounces (BASELINE configs[2])",
    "c4": "_TERRAIN + glass walls 128^3 chunk, 3840x2160 multi-bounce (BASELINE configs[3])",
    "c1": "_GLASS_CUBE + _HIGH_PERFORMANCE + _COLOR_ONLY 32^3, 400x400 (BASELINE configs[0])",
    "c5": "synthetic 1024^3 procedural chunk, 7680x4320, 64 orbit poses, maxRayLength 2048 (BASELINE configs[4])",
    "c5s": "synthetic 512^3 procedural chunk, 3840x2160, 64 orbit poses, maxRayLength 1024 (half-size rehearsal of configs[4])",
}
I_STEP = 16          # algorithmic thread-instructions per reference DDA step
NCU_SUMMARY = os.path.join(ROOT, "profiles", "ncu_summary.json")


def ncu_entry(workload):
    """Counters of the timed kernel from the committed `ncu --set full` capture of this workload (profiles/ncu_summary.json, written
    by tools/ncu_summary.py from the .ncu-rep; per-launch means over the captured launches, one per cycled pose).  None if the
    workload was never captured: nothing is typed into this file."""
    try:
        return json.load(open(NCU_SUMMARY)).get(workload)
    except Exception:
        return None


def config_of(args, poses, W, H):
    """The `config` object: the workload only, IDENTICAL in both arms (what differs between the arms is under `run`)."""
    return {"workload": WORKLOAD_DESC[args.workload], "poses": poses if len(poses) <= 8 else "%d orbit poses" % len(poses), "frame": [W, H],
            "maxRayLength": MAX_LEN.get(args.workload, 100.0),
            "inputs": "reference texture tiles (tests/golden/ref_tiles.npz) for c1-c4, procedural atlas for c5; seeded value-noise heightmap",
            "l2": "GPU arm: flushed by a 256 MiB fill before every timed step window (the fill itself is outside the window); CPU arm: n/a"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--gather", default=None, choices=["p2p", "dma", "nccl", "none"],
                    help="N>1 (default: p2p with --scaling strong, where the tiles of ONE frame have to meet; none with --scaling weak, where every GPU "
                         "renders its own frames - the fused gather of all frames into rank 0 is then measured in a second pass and reported as `gathered`): p2p = every rank's render kernel stores straight into rank 0's frame over NVLink (fused render+gather); "
                         "dma = every rank renders into its own memory and a copy-engine transfer on a side stream puts the pixels into rank 0's frame "
                         "(overlaps the next frame's render when frames are pipelined); "
                         "nccl = all_gather_into_tensor after the render; none = frames stay on their GPUs")
    ap.add_argument("--gather-format", default="rgb8", choices=["rgb8", "rgba8"],
                    help="N>1, p2p gather: pixel format of the frames rank 0 collects - packed RGB8 (the reference's colour attachment, FrameBuffer.cpp:8; "
                         "25 % fewer bytes over NVLink, default) or the ABI's RGBA8")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: one full frame per GPU per step (pose batch); strong: ONE frame per step, 16-row bands dealt round-robin")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--octants", type=int, default=1, choices=[0, 1], help="VRT_OPT_OCTANT_FIELDS (tuning comparison; 1 = default)")
    ap.add_argument("--radius-cap", type=int, default=0, help="VRT_OPT_RADIUS_CAP (tuning comparison; 0 = default)")
    ap.add_argument("--no-affinity", action="store_true", help="N > 1: do not restrict each rank to its GPU's local host cores")
    ap.add_argument("--e2e-depth", type=int, default=4, help="frames in flight in the overlapped e2e leg")
    ap.add_argument("--signal", default="flag", choices=["flag", "nccl"],
                    help="N>1, p2p gather: how rank 0 learns that a frame is complete - flag = every rank's stream adds to a counter in rank 0's memory after its "
                         "render (vrt_signal_add) and rank 0's stream waits on it (vrt_stream_wait_value): no collective, no host round trip; nccl = 4-byte all-reduce per step")
    ap.add_argument("--pipe-depth", type=int, default=4, help="frames in flight in the `pipelined` pass (and number of gather target buffers)")
    ap.add_argument("--remote-rows", type=int, default=1, choices=[0, 1], help="VRT_OPT_REMOTE_ROWS (tuning comparison; 1 = default)")
    ap.add_argument("--persistent", action="store_true", help="VRT_FLAG_PERSISTENT launch (tuning comparison)")
    return ap.parse_args()


# ------------------------------------------------------------------ the reference arm (CPU)
def row_blocks(height, rows):
    """The frame as consecutive blocks of `rows` scanlines (the last one shorter)."""
    return [(y, min(rows, height - y)) for y in range(0, height, rows)]


def cpu_render(common, p, grid, atlas, threads, use_ref):
    """One call into the reference shader (oracle/_ref, all `threads`; work items are 128-pixel row segments, so a block of a
    few dozen rows keeps every thread busy) - or the oracle port when _ref was never built."""
    if use_ref:
        common.ref_render(p, grid, atlas, nthreads=threads)
    else:
        common.oracle_render(p, grid, atlas, nthreads=threads)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import common
    common.build_oracle()
    sc, poses, W, H, flags = WORKLOADS[args.workload]
    grid, n, atlas, ts = common.scene(sc)
    threads = os.cpu_count() or 1
    use_ref = common.ref_lib() is not None
    frames = [mk(common, args.workload, W, H, sc, pose, flags) for pose in poses]
    # A step is a FULL frame when K + W of them fit in ~150 s on this box; otherwise a bounded sample: consecutive blocks of rows
    # (at least 4 rows per thread), step i taking block i of pose i, so that the steps sweep every frame from bottom to top.
    t0 = time.perf_counter()
    cpu_render(common, frames[0], grid, atlas, threads, use_ref)
    t_full = time.perf_counter() - t0
    budget = 150.0 / max(1, args.steps + args.warmup)
    rows = H
    if t_full > budget:
        rows = max(min(H, 4 * threads), int(H * budget / t_full))
    blocks = row_blocks(H, rows)

    def step_params(i):
        q = frames[i % len(frames)].copy()
        q.tile_y0, q.tile_h = blocks[i % len(blocks)]
        return q

    # rays of every distinct (pose, block) the timed steps touch, counted by the oracle outside the timing
    ray_cache = {}

    def rays_of(i):
        key = (i % len(frames), i % len(blocks))
        if key not in ray_cache:
            c = common.oracle_render(step_params(i), grid, atlas, nthreads=threads).counters
            ray_cache[key] = c["primary_rays"] + c["secondary_rays"] + c["shadow_rays"]
        return ray_cache[key]

    for i in range(args.warmup):
        cpu_render(common, step_params(i), grid, atlas, threads, use_ref)
    times = []
    total_r = 0
    for i in range(args.steps):
        q = step_params(i)
        t0 = time.perf_counter()
        cpu_render(common, q, grid, atlas, threads, use_ref)
        times.append(time.perf_counter() - t0)
    for i in range(args.steps):
        total_r += rays_of(i)
    total_t = sum(times)
    v = total_r / total_t / 1e6
    kind = "reference" if use_ref else "port"
    sample = ("full frames" if rows == H else "blocks of %d consecutive scanlines (1/%.1f of a frame per step), step i = block i of pose i" % (rows, H / rows)) + \
             "; rays counted by the oracle on the same pixels"
    line = {
        "impl": "reference", "metric": "Mrays/s at 1080p terrain scene" if args.workload == "c2" else "Mrays/s", "value": v, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / max(1, args.steps),
        "ms_per_step_median": 1e3 * float(np.median(times)), "ms_per_step_min": 1e3 * float(np.min(times)),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_of(args, poses, W, H),
        "run": {"traversal": "the reference's voxel.glsl compiled as C++ (oracle/_ref)" if use_ref else "oracle port of voxel.glsl",
                "sharding": "rank 0 only, %d host threads" % threads, "host_affinity": None},
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)
    return 0


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """Polls NVML (SM clock, throttle reasons, power) back to back from a thread while the timed region runs."""

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.stop_flag = False
        self.th = None
        self.err = None

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nv = nv
            self.h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_sm = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.th = threading.Thread(target=self._run, daemon=True)
            self.th.start()
        except Exception as e:       # no NVML: report it, never fake a clock
            self.err = repr(e)

    def _run(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM), nv.nvmlDeviceGetCurrentClocksEventReasons(self.h),
                                  nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0))
            except Exception as e:
                self.err = repr(e)
                return
            time.sleep(0.0003)

    def mark(self):
        return len(self.rows)

    def stop(self, first=0, last=None):
        self.stop_flag = True
        if self.th:
            self.th.join(timeout=2)
        rows = self.rows[first:last] or self.rows
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: %s" % self.err]}
        nv = self.nv
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        reasons = sorted(k for k, b in bits.items() if any(r[1] & b for r in rows))
        return {"sm_mhz": float(np.median([r[0] for r in rows])), "sm_max_mhz": float(self.max_sm), "samples": len(rows),
                "power_w_max": max(r[2] for r in rows), "reasons": reasons, "how": "NVML polled back to back (~0.4 ms per sample) during the timed region"}


# ------------------------------------------------------------------ our arm
def nvml_index(torch, local):
    """NVML index of CUDA device `local` (CUDA_VISIBLE_DEVICES may renumber)."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        want = "GPU-" + str(torch.cuda.get_device_properties(local).uuid)
        for i in range(nv.nvmlDeviceGetCount()):
            u = nv.nvmlDeviceGetUUID(nv.nvmlDeviceGetHandleByIndex(i))
            if (u.decode() if isinstance(u, bytes) else u) == want:
                return i
    except Exception:
        pass
    return local


def pin_to_gpu_cores(index):
    """Restricts this process to the host cores NVML reports as local to GPU `index` (NUMA placement of the pinned frame
    buffers follows the allocating thread).  Returns the number of cores, or None if it could not be done."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        words = (max(os.cpu_count() or 64, 64) + 63) // 64
        mask = nv.nvmlDeviceGetCpuAffinity(nv.nvmlDeviceGetHandleByIndex(index), words)
        cpus = {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1} & os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


def run_ours(args):
    import torch
    import torch.distributed as dist
    import common
    import voxelraytracer_b200 as vrt
    from voxelraytracer_b200 import sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    # N > 1: each rank's frames go to pinned host memory next to ITS GPU (8 ranks allocating on one socket halve the e2e leg)
    host_cores = pin_to_gpu_cores(nvml_index(torch, local)) if world > 1 and not args.no_affinity else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc, poses, W, H, flags = WORKLOADS[args.workload]
    grid, n, atlas, ts = common.scene(sc)
    ctx = vrt.Context(local)
    ctx.set_option(vrt.OPT_OCTANT_FIELDS, args.octants)
    ctx.set_option(vrt.OPT_RADIUS_CAP, args.radius_cap)
    ctx.set_option(vrt.OPT_REMOTE_ROWS, args.remote_rows)
    t_up0 = time.perf_counter()
    ctx.upload_grid(grid, n)
    t_upload = time.perf_counter() - t_up0
    ctx.upload_atlas(atlas, ts)
    if args.mode == "exact":
        flags |= vrt.FLAG_EXACT
    if args.persistent:
        flags |= vrt.FLAG_PERSISTENT
    strong = args.scaling == "strong" and world > 1
    gather = "none" if world == 1 else (args.gather or ("p2p" if strong else "none"))
    also_gathered = world > 1 and not strong and gather == "none"                   # weak scaling: second pass with the fused gather
    BR = 16
    NBUF = max(2, args.pipe_depth)                                                   # target buffers (frames that may be in flight)
    npix = W * H
    ipc = gather in ("p2p", "dma")                                                # rank 0 owns the gathered frames, the others map them
    want_ipc = ipc or also_gathered
    gbpp = 3 if (want_ipc and args.gather_format == "rgb8") else 4                  # bytes per pixel of the gathered frames
    frame_bytes = npix * gbpp
    if strong:      # ONE frame per step: every rank renders its 16-row bands of the same pose
        full = [mk(common, args.workload, W, H, sc, poses[i % len(poses)], flags) for i in range(len(poses))]
        frames = [sharding.band_params(p, rank, world, BR) for p in full]
    else:           # one full frame per rank per step, each rank on its own pose sequence
        frames = [mk(common, args.workload, W, H, sc, poses[(i + rank) % len(poses)], flags) for i in range(len(poses))]
    frames = frames[:max(1, min(len(frames), args.steps))]      # only the poses the timed steps will cycle through
    my_pix = vrt.tile_pixels(frames[0])

    def gathered_form(p):
        """The params of a frame as it is rendered INTO the gather target (packed RGB8 when that is the gather format)."""
        if gbpp == 4:
            return p
        q = p.copy()
        q.flags |= vrt.FLAG_RGB8_OUT
        return q
    gframes = [gathered_form(p) for p in frames]

    # reference-equivalent ray / step counts of what THIS rank renders (exact kernel's counters == oracle's, see tests)
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

    stream = torch.cuda.Stream()
    comm = torch.cuda.Stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")                        # > 126 MB L2
    token = torch.zeros(1, dtype=torch.int32, device="cuda")
    slots = 1 if strong else world                                                          # frames held by rank 0 per step
    local_fb = [torch.empty(npix * 4, dtype=torch.uint8, device="cuda") for _ in range(NBUF)]
    gathered = None
    p2p_base = None
    if want_ipc:
        # rank 0 owns the gathered frames (double-buffered); the others map them through CUDA IPC and store into them.
        # If the mapping fails on any rank (no peer access between the devices), every rank falls back to the NCCL gather.
        ok = 1
        try:
            if rank == 0:
                p2p_base, handle = ctx.ipc_alloc_frame(NBUF * slots * frame_bytes + 8192)
                box = [handle]
            else:
                box = [None]
        except vrt.VrtError as e:
            print("bench: ipc_alloc_frame failed: %s" % e, file=sys.stderr)
            box, ok = [None], 0
        dist.broadcast_object_list(box, src=0)
        if rank != 0 and box[0] is not None:
            try:
                p2p_base = ctx.ipc_open(box[0])
            except vrt.VrtError as e:
                print("bench: ipc_open failed on rank %d: %s" % (rank, e), file=sys.stderr)
                ok = 0
        if box[0] is None:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            p2p_base = int(p2p_base.value)
            cnt_base = p2p_base + ((NBUF * slots * frame_bytes + 4095) // 4096) * 4096      # completion counters, one per target buffer
            if rank == 0:
                ctx.device_memset(cnt_base, 0, 4096); ctx.synchronize()
            dist.barrier()
        else:
            gather, p2p_base, ipc, want_ipc, also_gathered = ("nccl" if ipc else gather), None, False, False, False
    # What the gather can reach at best: every rank but 0 copies a frame-sized buffer into its slot of rank 0's memory, all at once
    # (plain device-to-peer copy kernels, no rendering): aggregate GB/s into rank 0 = the NVLink ingest ceiling of this box.
    ingest_peak_gbs = None
    if want_ipc:
        reps = 20
        src = torch.empty(frame_bytes, dtype=torch.uint8, device="cuda")
        dst = _wrap_device(torch, p2p_base + (rank if not strong else 0) * frame_bytes, frame_bytes) if rank != 0 else None
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if rank != 0:
            for _ in range(reps):
                dst.copy_(src)
        e1.record(); torch.cuda.synchronize()
        tt = torch.tensor([e0.elapsed_time(e1) / 1e3 if rank != 0 else 0.0], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ingest_peak_gbs = (world - 1) * frame_bytes * reps / float(tt.item()) / 1e9
        dist.barrier()
    if gather == "nccl":
        per = sharding.max_rows(H, world, BR) * W * 4 if strong else frame_bytes          # nccl gather: always RGBA8
        gathered = torch.empty(world * per, dtype=torch.uint8, device="cuda")
        local_fb = [torch.zeros(per, dtype=torch.uint8, device="cuda") for _ in range(NBUF)]

    def target(i):
        """(params, device pointer) of step i's output for this rank."""
        p = gframes[i % len(frames)] if ipc else frames[i % len(frames)]
        if gather == "p2p":
            if strong:
                q = p.copy()
                q.flags |= vrt.FLAG_OUT_FRAME          # bands land at their frame rows inside rank 0's frame
                return q, p2p_base + (i % NBUF) * frame_bytes
            return p, p2p_base + ((i % NBUF) * world + rank) * frame_bytes
        return p, local_fb[i % NBUF].data_ptr()

    uses = [0] * NBUF
    sig_flag = args.signal == "flag"

    def completion(i, st_render, rendered):
        """p2p gather: what tells rank 0 that every rank's stores of step i have landed.  Returns the event a consumer would wait on."""
        slot = i % NBUF
        uses[slot] += 1
        done = torch.cuda.Event()
        if sig_flag:
            ctx.signal_add(cnt_base + 4 * slot, 1, st_render.cuda_stream)           # after this rank's render, on its stream
            if rank == 0:
                comm.wait_event(rendered)
                ctx.stream_wait_value(cnt_base + 4 * slot, uses[slot] * world, comm.cuda_stream)
                done.record(comm)
            else:
                done.record(st_render)
        else:
            with torch.cuda.stream(comm):
                comm.wait_event(rendered)
                dist.all_reduce(token)               # 4-byte all-reduce on the side stream
                done.record(comm)
        return done

    def dma_gather(i, st):
        """gather == dma: this rank's pixels of step i, rendered compactly into local_fb, go to rank 0's frame through the copy engines."""
        src = local_fb[i % NBUF].data_ptr()
        if strong:
            chunk = BR * W * gbpp                                      # one band of rows, contiguous on both sides
            rows_mine = my_pix // W
            nb, tail = rows_mine // BR, rows_mine % BR
            dst = p2p_base + (i % NBUF) * frame_bytes + rank * chunk
            if nb:
                ctx.copy_2d_async(dst, world * chunk, src, chunk, chunk, nb, st.cuda_stream)
            if tail:                                                    # a last, shorter band
                ctx.copy_2d_async(dst + nb * world * chunk, world * chunk, src + nb * chunk, chunk, tail * W * gbpp, 1, st.cuda_stream)
        else:
            ctx.copy_2d_async(p2p_base + ((i % NBUF) * world + rank) * frame_bytes, frame_bytes, src, frame_bytes, frame_bytes, 1, st.cuda_stream)

    def one_step(i, ev0=None, ev1=None, do_flush=True):
        """flush (untimed) -> [ev0] render (p2p: its stores ARE the gather) (+ nccl gather) [ev1].  Returns the completion event."""
        p, ptr = target(i)
        with torch.cuda.stream(stream):
            if do_flush:
                flush.fill_(i & 0xFF)                # evict L2 between timed windows (the window starts after it)
            if ev0 is not None:
                ev0.record(stream)
            ctx.render_to(p, ptr, stream.cuda_stream)
            if ev1 is not None and gather not in ("nccl", "dma"):
                ev1.record(stream)
        done = torch.cuda.Event()
        if gather == "none":
            done.record(stream)
            return done
        rendered = torch.cuda.Event()
        rendered.record(stream)
        if gather == "p2p":
            return completion(i, stream, rendered)
        with torch.cuda.stream(comm):
            comm.wait_event(rendered)
            if gather == "nccl":
                dist.all_gather_into_tensor(gathered, local_fb[i % NBUF])
                if ev1 is not None:
                    ev1.record(comm)
            elif gather == "dma":
                dma_gather(i, comm)
                if ev1 is not None:
                    ev1.record(comm)                 # the window includes this rank's transfer
                dist.all_reduce(token)
            else:
                # p2p: 4-byte all-reduce = "every rank's stores of step i have landed in rank 0's frame"; it runs on the
                # comm stream beside the next step's flush + render and is what rank 0's consumer would wait on
                dist.all_reduce(token)
            done.record(comm)
        return done

    # ---- gather check (untimed): what rank 0 holds after one step equals what the ranks rendered locally
    gather_check = None
    if gather != "none":
        one_step(0, do_flush=False).synchronize()
        torch.cuda.synchronize(); dist.barrier()
        loc = torch.empty(frame_bytes, dtype=torch.uint8, device="cuda")
        ctx.render_to(gathered_form(full[0]) if strong else gframes[0], loc.data_ptr(), None)      # strong: rank-local render of the WHOLE frame
        torch.cuda.synchronize()
        held = _wrap_device(torch, p2p_base, slots * frame_bytes) if (ipc and rank == 0) else gathered
        if strong:
            if rank == 0:
                if gather == "nccl":
                    mr = sharding.max_rows(H, world, BR)
                    perm = torch.from_numpy(sharding.row_permutation(H, world, BR)).cuda()
                    held = held.view(world * mr, W * 4)[perm].reshape(-1)
                gather_check = "ok" if torch.equal(held[:frame_bytes], loc) else "MISMATCH"
        else:
            mine = loc.to(torch.int64).sum().reshape(1)
            sums = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(sums, mine)
            if rank == 0:
                ok = torch.equal(held[:frame_bytes], loc)
                for r in range(world):
                    ok = ok and int(held[r * frame_bytes:(r + 1) * frame_bytes].to(torch.int64).sum()) == int(sums[r])
                gather_check = "ok" if ok else "MISMATCH"
        dist.barrier()

    sampler = ClockSampler(nvml_index(torch, local))
    if rank == 0:
        sampler.start()
    for i in range(args.warmup):
        stream.wait_event(one_step(i))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    # Timed region: K steps, each its own CUDA-event window on the render stream, L2 flushed before each window.
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = ctx.kernel_launches()
    m0 = sampler.mark()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        stream.wait_event(one_step(i, ev[i][0], ev[i][1]))     # step i+1 starts after step i (and its gather / signal) finished
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t_wall0
    m1 = sampler.mark()
    launches_timed = ctx.kernel_launches() - launches0
    step_ms = [a.elapsed_time(b) for a, b in ev]
    t_steps = sum(step_ms) / 1e3
    if os.environ.get("BENCH_DUMP_STEPS"):            # per-rank step times for offline analysis (tools/)
        np.save(os.path.join(os.environ["BENCH_DUMP_STEPS"], "steps_rank%d.npy" % rank), np.array(step_ms))

    # kernel-only windows (for the roofline): the same steps without a gather inside the window
    if gather in ("nccl", "dma"):
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for i in range(args.steps):
            p, ptr = target(i)
            with torch.cuda.stream(stream):
                flush.fill_(i & 0xFF)
                kev[i][0].record(stream)
                ctx.render_to(p, ptr, stream.cuda_stream)
                kev[i][1].record(stream)
        torch.cuda.synchronize()
        t_kernel = sum(a.elapsed_time(b) for a, b in kev) / 1e3
    else:
        t_kernel = t_steps
    # The kernel's own duration, for the roofline: the same K steps (same L2 flush before each), each read back through the library's
    # event pair, which is recorded directly around the launch on the launching stream (vrt_last_render_ms); the step windows above also
    # hold torch's event records and the gap between them and the launch.
    kms = []
    for i in range(args.steps):
        p, ptr = target(i)
        with torch.cuda.stream(stream):
            flush.fill_(i & 0xFF)
            ctx.render_to(p, ptr, stream.cuda_stream)
        kms.append(ctx.last_render_ms())
    t_kernel = sum(kms) / 1e3
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop(m0, m1) if rank == 0 else None
    rays = sum(counts[i % len(frames)]["rays"] for i in range(args.steps))
    steps_ref = sum(counts[i % len(frames)]["steps"] for i in range(args.steps))
    steps_act = sum(counts[i % len(frames)]["actual"] for i in range(args.steps))
    hits = sum(counts[i % len(frames)]["hits"] for i in range(args.steps))

    # Pipelined pass (informational, not the headline): the same K steps with two frames in flight - renders alternate
    # between two streams (double-buffered targets), no L2 fills - so one frame's tail overlaps the next frame's ramp-up
    # and, for N > 1, its gather / completion signal.  One CUDA-event pair around the whole region, max over ranks.
    t_pipeline = None
    if True:
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        streams2 = [stream] + [torch.cuda.Stream() for _ in range(NBUF - 1)]
        last = [None] * NBUF

        def pipe_step(i):
            p, ptr = target(i)
            st = streams2[i % NBUF]
            with torch.cuda.stream(st):
                if last[i % NBUF] is not None:
                    st.wait_event(last[i % NBUF])        # buffer i % NBUF is free once step i - NBUF (and its gather) finished
                ctx.render_to(p, ptr, st.cuda_stream)
                rendered = torch.cuda.Event(); rendered.record(st)
            done = rendered
            if gather == "p2p":
                done = completion(i, st, rendered)
            elif gather != "none":
                with torch.cuda.stream(comm):
                    comm.wait_event(rendered)
                    if gather == "nccl":
                        dist.all_gather_into_tensor(gathered, local_fb[i % NBUF])
                    elif gather == "dma":
                        dma_gather(i, comm)
                        dist.all_reduce(token)
                    else:
                        dist.all_reduce(token)
                    done = torch.cuda.Event(); done.record(comm)
            last[i % NBUF] = done

        for i in range(2 * NBUF):
            pipe_step(i)
        torch.cuda.synchronize()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record(stream)
        for st_ in streams2[1:]:
            st_.wait_event(p0)
        for i in range(args.steps):
            pipe_step(i)
        for e in last:
            stream.wait_event(e)
        for st_ in streams2[1:]:
            stream.wait_stream(st_)
        p1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t_pipeline = p0.elapsed_time(p1) / 1e3

    # Weak scaling, second pass: the same K steps with every rank's render kernel storing its frame straight into rank 0's memory over
    # NVLink (the fused render + gather of the tile-sharded mode applied to whole frames): what collecting ALL frames on one GPU costs.
    gathered_pass = None
    if also_gathered:
        gev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        torch.cuda.synchronize(); dist.barrier()
        for i in range(-3, args.steps):
            with torch.cuda.stream(stream):
                flush.fill_(i & 0xFF)
                if i >= 0:
                    gev[i][0].record(stream)
                ctx.render_to(gframes[i % len(frames)], p2p_base + ((i % NBUF) * world + rank) * frame_bytes, stream.cuda_stream)
                if i >= 0:
                    gev[i][1].record(stream)
                rendered = torch.cuda.Event(); rendered.record(stream)
            done = completion(i + 3 * NBUF, stream, rendered)      # "all stores of this step have landed" (i starts at -3; slots stay aligned)
            stream.wait_event(done)
        torch.cuda.synchronize(); dist.barrier()
        tg = torch.tensor([sum(a.elapsed_time(b) for a, b in gev) / 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(tg, op=dist.ReduceOp.MAX)
        # check: rank 0 holds every rank's last frame (sums, like gather_check)
        loc = torch.empty(frame_bytes, dtype=torch.uint8, device="cuda")
        last = args.steps - 1
        ctx.render_to(gframes[last % len(frames)], loc.data_ptr(), None); torch.cuda.synchronize()
        mine = loc.to(torch.int64).sum().reshape(1)
        sums = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(sums, mine)
        ok = None
        if rank == 0:
            held = _wrap_device(torch, p2p_base + (last % NBUF) * world * frame_bytes, world * frame_bytes)
            ok = all(int(held[r * frame_bytes:(r + 1) * frame_bytes].to(torch.int64).sum()) == int(sums[r]) for r in range(world))
        gathered_pass = {"t": float(tg.item()), "ok": ok}
        dist.barrier()

    # e2e through the host-facing API, every step: params by value in (h2d), this rank's pixels back in pinned host memory (d2h).
    #  sync form : vrt_render_host - render, copy, synchronise, one frame at a time
    #  overlapped: vrt_render_to + vrt_copy_to_host_async on two streams with two device/host buffers (two frames in flight), the
    #              double-buffered form a frame loop would use; every frame still lands in host memory inside the timed region
    host, hp = ctx.host_alloc(my_pix * 4)

    def sync_leg(direct):
        ctx.set_option(vrt.OPT_HOST_DIRECT, direct)
        for i in range(3):
            ctx.render_host(frames[i % len(frames)], hp, stream.cuda_stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            ctx.render_host(frames[i % len(frames)], hp, stream.cuda_stream)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    t_sync_staged = sync_leg(0)                      # device frame, then a copy
    t_e2e_sync = sync_leg(1)                         # the kernel stores into the pinned buffer (VRT_OPT_HOST_DIRECT, the default)
    D = max(2, args.e2e_depth)                       # frames in flight
    hosts = [ctx.host_alloc(my_pix * 4) for _ in range(D)]
    dbuf = [torch.empty(my_pix * 4, dtype=torch.uint8, device="cuda") for _ in range(D)]
    st2 = [torch.cuda.Stream() for _ in range(D)]

    def e2e_step(i):
        st = st2[i % D]
        ctx.render_to(frames[i % len(frames)], dbuf[i % D].data_ptr(), st.cuda_stream)
        ctx.copy_to_host_async(hosts[i % D][1], dbuf[i % D].data_ptr(), my_pix * 4, st.cuda_stream)

    def overlapped(step_fn):
        for i in range(2 * D):
            step_fn(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            if i >= D:
                st2[i % D].synchronize()             # frame i-D is in host memory (a consumer would read it here)
            step_fn(i)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    def e2e_direct(i):                               # no device frame, no copy: the render kernel's row stores cross PCIe
        ctx.render_to(frames[i % len(frames)], hosts[i % D][1], st2[i % D].cuda_stream)

    frames_rgb = []
    for f in frames:
        g = f.copy(); g.flags |= vrt.FLAG_RGB8_OUT; frames_rgb.append(g)

    def e2e_direct_rgb(i):                           # same, packed RGB8 (the reference's attachment format): 25 % fewer bytes over the link
        ctx.render_to(frames_rgb[i % len(frames)], hosts[i % D][1], st2[i % D].cuda_stream)

    t_e2e_rgb = overlapped(e2e_direct_rgb)
    rgb_ok = bool(np.array_equal(hosts[(args.steps - 1) % D][0][:my_pix * 3].reshape(-1, 3),
                                 ctx_frame_host(ctx, frames[(args.steps - 1) % len(frames)], my_pix).reshape(-1, 4)[:, :3]))
    t_e2e_direct = overlapped(e2e_direct)
    direct_ok = bool(np.array_equal(hosts[(args.steps - 1) % D][0], ctx_frame_host(ctx, frames[(args.steps - 1) % len(frames)], my_pix)))
    t_e2e_staged = overlapped(e2e_step)
    e2e_ok = bool(np.array_equal(hosts[(args.steps - 1) % D][0], dbuf[(args.steps - 1) % D].cpu().numpy())) and direct_ok
    # the bound of the overlapped form: the same loop with the render left out (device -> pinned host copies only)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()                               # all ranks copy at the same time: the box's host-ingest ceiling
    t0 = time.perf_counter()
    for i in range(args.steps):
        if i >= D:
            st2[i % D].synchronize()
        ctx.copy_to_host_async(hosts[i % D][1], dbuf[i % D].data_ptr(), my_pix * 4, st2[i % D].cuda_stream)
    torch.cuda.synchronize()
    t_copy = time.perf_counter() - t0

    d2h_all = my_pix * 4 * args.steps / t_copy / 1e9
    if world > 1:      # the host-ingest ceiling of the box: every rank's copies-only leg ran at the same time
        dd = torch.tensor([d2h_all], dtype=torch.float64, device="cuda")
        dist.all_reduce(dd, op=dist.ReduceOp.SUM)
        d2h_all = float(dd.item())
    # max over ranks for times, sum over ranks for work
    if world > 1:
        v = torch.tensor([t_steps, t_kernel, t_e2e_direct, t_pipeline or 0.0, t_e2e_sync, t_e2e_staged, t_sync_staged], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        t_step_total, t_kernel_max, t_e2e_direct, t_pipeline, t_e2e_sync, t_e2e_staged, t_sync_staged = [float(x) for x in v.cpu()]
        s = torch.tensor([rays, steps_ref, steps_act, hits], dtype=torch.float64, device="cuda")
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        rays_all, steps_all, act_all, hits_all = [float(x) for x in s.cpu()]
    else:
        t_step_total, t_kernel_max = t_steps, t_kernel
        rays_all, steps_all, act_all, hits_all = float(rays), float(steps_ref), float(steps_act), float(hits)

    t_e2e = min(t_e2e_direct, t_e2e_staged)
    l2_gbs = hbm_probe_gbs = None
    if rank == 0:
        # denominators measured on this box, outside every timed region: L2-resident and HBM-sized read sweeps
        l2_gbs = ctx.probe_read_gbs(32 << 20, 400)
        hbm_probe_gbs = ctx.probe_read_gbs(4 << 30, 4)
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
        # dominant kernel = render_kernel; per launch (rank 0's launches): algorithmic instructions and bytes
        per_launch_s = t_kernel / args.steps
        instr_per_launch = I_STEP * steps_ref / args.steps
        bytes_per_launch = (steps_ref * 1 + hits * 4 + my_pix * 4 * args.steps) / args.steps
        ach_issue = instr_per_launch / per_launch_s
        frames_total = args.steps * (1 if strong else world)
        shard = "one GPU" if world == 1 else (
            ("ONE frame per step split into 16-row bands dealt round-robin over the ranks" if strong else
             "one frame per GPU per step (pose batch): independent units, no data-path collective - every frame stays in (and, in the e2e leg, goes to the host from) the GPU that rendered it")
            + ", grid replicated; gather: " + {"p2p": "every rank's render kernel stores its pixels straight into rank 0's frame over NVLink (CUDA IPC mapping), "
                                                      "completion signal per step: " + ("a counter in rank 0's memory that every rank's stream adds to after its render and rank 0's stream waits on (no collective)" if args.signal == "flag" else "4-byte NCCL all-reduce (overlapped)"),
                                               "dma": "every rank renders into its own memory; a copy-engine transfer on a side stream (cudaMemcpy2DAsync into rank 0's frame, CUDA IPC mapping) "
                                                      "follows inside the step window and overlaps the next frame's render in the pipelined pass",
                                               "nccl": "all_gather_into_tensor after the render, inside the step window",
                                               "none": "none in the timed steps; the fused NVLink gather of all frames into rank 0 is measured in a second pass (`gathered`)"}[gather])
        # What the timed kernel really executes, from the committed ncu capture of this workload at 1 GPU (per-launch means over the
        # cycled poses); the TIME it is divided by is this run's own CUDA-event time.  Only meaningful when the launches are the
        # captured ones: one GPU, full frames, default traversal.
        nc = ncu_entry(args.workload) if (world == 1 and args.mode == "fast" and not args.persistent) else None
        exec_per_launch = nc["thread_inst_per_launch"] if nc else None
        ach_actual = exec_per_launch / per_launch_s if nc else None
        traffic = nc["dram_bytes_per_launch"] if nc else None
        roof = {"bound": "issue", "unit": "Tinstr/s", "peak": p_issue / 1e12,
                "achieved": ach_actual / 1e12 if nc else None, "frac": ach_actual / p_issue if nc else None,
                "frac_actual": ach_actual / p_issue if nc else None,
                "frac_frames_in_flight": (exec_per_launch / (t_pipeline / args.steps) / p_issue) if (nc and t_pipeline) else None,
                "frac_frames_in_flight_def": "same executed thread-instructions over the time per frame of the `pipelined` pass (frames in flight on several streams, "
                                             "no L2 fills): the launch latency, ramp-up and tail of one launch overlap the next, so this is what the kernel sustains; "
                                             "`frac` is one isolated, L2-flushed launch",
                "achieved_def": "thread-instructions the kernel EXECUTES per launch (ncu smsp__inst_executed x lanes per instruction, mean over the cycled poses) "
                                "/ this run's CUDA-event time per launch",
                "peak_source": "%d SM x 4 schedulers x 32 lanes x sampled SM clock %.0f MHz" % (sms, f_mhz),
                "traffic": traffic, "traffic_unit": "DRAM bytes read + written per launch (ncu --set full)",
                "effective": {"achieved": ach_issue / 1e12, "frac": ach_issue / p_issue,
                              "def": "SURVEY.md 8(d) algorithmic figure: %d thread-instr per REFERENCE DDA step x reference steps / time; empty-space skipping "
                                     "executes far fewer steps than the reference takes, so this is a speed-up over the step floor, not a utilisation" % I_STEP},
                "ncu": None if not nc else {k: nc[k] for k in ("file", "kernel", "commit", "launches", "duration_us_per_launch", "issue_active_pct",
                                                                "lanes_per_inst", "warp_inst_per_launch", "thread_inst_per_pixel") if k in nc}}
        dram_gbs = traffic / per_launch_s / 1e9 if nc else None
        line = {
            "metric": "Mrays/s at 1080p terrain scene" if args.workload == "c2" else "Mrays/s", "value": rays_all / t_step_total / 1e6,
            "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_step_total / args.steps,
            "ms_per_step_median": float(np.median(step_ms)), "ms_per_step_min": float(np.min(step_ms)),
            "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(args, poses, W, H),
            "run": {"traversal": args.mode, "sharding": shard,
                    "host_affinity": None if world == 1 else ("rank 0 on the %d host cores local to its GPU (NVML), likewise every rank" % host_cores
                                                              if host_cores else "not set")},
            "frames_per_s": frames_total / t_step_total,
            "kernel_ms_per_launch": 1e3 * t_kernel_max / args.steps, "kernel_ms_per_launch_median": float(np.median(kms)), "kernel_ms_per_launch_min": float(np.min(kms)),
            "kernel_ms_def": "render kernel only: the library's CUDA-event pair around the launch on the launching stream (vrt_last_render_ms), separate pass of the same K steps, L2 flushed before each",
            "rays_per_frame": rays_all / frames_total, "ref_steps_per_frame": steps_all / frames_total,
            "actual_steps_per_frame": act_all / frames_total,
            "gsteps_per_s": steps_all / t_step_total / 1e9,
            "roofline": roof,
            "roofline_hbm": {"bound": "hbm", "achieved": dram_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": dram_gbs / hbm_peak if nc else None,
                             "achieved_def": "measured DRAM traffic per launch (ncu) / this run's time per launch: HBM is not the bound of this path",
                             "algorithmic_gbs": bytes_per_launch / per_launch_s / 1e9,
                             "algorithmic_def": "1 B per reference DDA step + 4 B per shaded hit + 4 B per pixel (SURVEY.md 8(d)); mostly bytes the skipping walk never touches",
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "6650 GB/s (of fallback)"},
            "roofline_l2": {"bound": "l2", "achieved": bytes_per_launch / per_launch_s / 1e9, "peak": l2_gbs, "unit": "GB/s",
                            "frac": bytes_per_launch / per_launch_s / 1e9 / l2_gbs if l2_gbs else None,
                            "achieved_def": "algorithmic bytes (see roofline_hbm.algorithmic_def) / time",
                            "peak_source": "vrt_probe_read_gbs: 32 MiB buffer read 400 times by one kernel on this GPU, just now",
                            "hbm_probe_gbs": hbm_probe_gbs, "hbm_probe": "same kernel, 4 GiB buffer, 4 passes"},
            "e2e": {"value": rays_all / t_e2e / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": C.sizeof(vrt.Params), "d2h_bytes_per_step": my_pix * 4,
                    "call": ("vrt_render_to into pinned host memory (the render kernel's 128-byte row stores cross PCIe; no device frame, no copy)"
                             if t_e2e_direct <= t_e2e_staged else "vrt_render_to + vrt_copy_to_host_async")
                            + ", %d frames in flight on %d streams (params by value, RGBA8 pixels in pinned host memory every step), per rank" % (D, D),
                    "direct_value": rays_all / t_e2e_direct / 1e6, "staged_value": rays_all / t_e2e_staged / 1e6,
                    "staged_call": "vrt_render_to (device frame) + vrt_copy_to_host_async, same streams",
                    "sync_value": rays_all / t_e2e_sync / 1e6,
                    "sync_call": "vrt_render_host into pinned memory (kernel stores straight to the host), synchronised, one frame at a time",
                    "sync_staged_value": rays_all / t_sync_staged / 1e6, "sync_staged_call": "vrt_render_host with VRT_OPT_HOST_DIRECT = 0: render, copy, synchronise",
                    "rgb8_value": rays_all / t_e2e_rgb / 1e6 if world == 1 else None,
                    "rgb8_call": "vrt_render_to into pinned host memory with VRT_FLAG_RGB8_OUT (%d B per step; not the headline: the ABI's default frame is RGBA8)" % (my_pix * 3),
                    "host_frame_check": e2e_ok and rgb_ok,
                    "d2h_only_ms_per_step": 1e3 * t_copy / args.steps, "d2h_only_gbs": my_pix * 4 * args.steps / t_copy / 1e9,
                    "d2h_only_gbs_all_ranks": d2h_all,
                    "frac_of_d2h_bound": t_copy / t_e2e if world == 1 else None},
            "gpu_launches": int(launches_timed),      # render_kernel launches of rank 0 inside the timed region (one per step)
            "gather_check": gather_check,
            "gathered": None if not gathered_pass else {
                "value": rays_all / gathered_pass["t"] / 1e6, "unit": "Mrays/s", "ms_per_step": 1e3 * gathered_pass["t"] / args.steps,
                "what": "second pass of the same K steps: every rank's render kernel stores its frame straight into rank 0's memory over NVLink (fused "
                        "render + gather, %s), completion signal per step (--signal %s); bound by what ONE GPU can ingest" % ("packed RGB8" if gbpp == 3 else "RGBA8", args.signal),
                "bytes_into_rank0_per_step": (world - 1) * frame_bytes, "achieved_gbs": (world - 1) * frame_bytes * args.steps / gathered_pass["t"] / 1e9,
                "ingest_peak_gbs": ingest_peak_gbs, "check": "ok" if gathered_pass["ok"] else "MISMATCH"},
            "gather": None if not ipc else {
                "how": "render kernel's own stores (fused)" if gather == "p2p" else "copy engines after a render into local memory (dma)",
                "format": "packed RGB8 (the reference's colour attachment format)" if gbpp == 3 else "RGBA8",
                "bytes_into_rank0_per_step": (world - 1) * frame_bytes if not strong else frame_bytes * (world - 1) // world,
                "achieved_gbs": ((world - 1) * frame_bytes if not strong else frame_bytes * (world - 1) / world) * args.steps / t_step_total / 1e9,
                "ingest_peak_gbs": ingest_peak_gbs,
                "ingest_peak_def": "all other ranks copying frame-sized buffers into rank 0's memory at once (device-to-peer copy kernels, no rendering), measured in this run",
                "pipelined_gbs": None if not t_pipeline else ((world - 1) * frame_bytes if not strong else frame_bytes * (world - 1) / world) * args.steps / t_pipeline / 1e9},
            "upload_grid_s": t_upload, "radius_fields": "8 direction-octant fields" if args.octants else "1 isotropic field",
            "pipelined": None if not t_pipeline else {"value": rays_all / t_pipeline / 1e6, "unit": "Mrays/s",
                                                       "frames_per_s": frames_total / t_pipeline,
                                                       "note": "%d frames in flight on %d streams (as many target buffers), no L2 fills, one event pair around all K steps" % (NBUF, NBUF)},
            "clocks": clocks,
            "wall_s": t_wall,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(common, frames, poses, grid, atlas)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()
    return 0


def ctx_frame_host(ctx, p, npix):
    """The same frame through vrt_render + vrt_read_framebuffer (device frame, blocking copy), as the check for the direct leg."""
    ctx.render(p)
    ctx.synchronize()
    return ctx.read_framebuffer(p).reshape(-1)[:npix * 4]


def _wrap_device(torch, ptr, nbytes):
    """A torch uint8 view of raw device memory (used to look at the IPC frame rank 0 owns)."""
    class _Raw:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 3}
    return torch.as_tensor(_Raw(), device="cuda")


def cpu_baseline(common, frames, poses, grid, atlas, target_s=12.0):
    """The reference shader (oracle/_ref) on all host cores on a BOUNDED sample of the workload: full frames of the recorded poses when
    a frame takes under ~2 s, otherwise blocks of consecutive scanlines sized for ~1 s each (step i = block i of pose i, so the sample
    sweeps the frames bottom to top), until ~target_s of wall time has been spent.  Rays are counted by the oracle on the same pixels,
    outside the timing - and only for the blocks that were timed (an 8K frame of the 1024^3 chunk is minutes of oracle time)."""
    threads = os.cpu_count() or 1
    use_ref = common.ref_lib() is not None
    H = frames[0].height
    probe = frames[0].copy()                      # how long does a scanline take?  time a thin block in the middle of the frame
    probe.tile_y0, probe.tile_h = H // 2, min(H, max(8, threads // 2))
    t0 = time.perf_counter()
    cpu_render(common, probe, grid, atlas, threads, use_ref)
    per_row = (time.perf_counter() - t0) / probe.tile_h
    rows = H if per_row * H <= 2.0 else max(8, min(H, int(1.0 / per_row)))
    blocks = row_blocks(H, rows)
    t_total, n, timed = 0.0, 0, []
    while t_total < target_s and n < 64 and (n < len(frames) or rows == H or t_total < target_s):
        q = frames[n % len(frames)].copy()
        q.tile_y0, q.tile_h = blocks[(n * 7) % len(blocks)]          # stride through the frame: sky, horizon and ground all get sampled
        t0 = time.perf_counter()
        cpu_render(common, q, grid, atlas, threads, use_ref)
        t_total += time.perf_counter() - t0
        timed.append(q)
        n += 1
        if rows == H and n >= len(frames) and t_total >= target_s:
            break
    r_total = 0
    for q in timed:
        c = common.oracle_render(q, grid, atlas, nthreads=threads).counters
        r_total += c["primary_rays"] + c["secondary_rays"] + c["shadow_rays"]
    what = "%d full frames" % n if rows == H else "%d blocks of %d consecutive scanlines (1/%.0f of a frame each)" % (n, rows, H / rows)
    return {"value": r_total / t_total / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "reference" if use_ref else "port",
            "sample": "%s (poses %s cycled), %.1f s of wall time on %d threads" % (what, ",".join(poses[:4]) + ("..." if len(poses) > 4 else ""), t_total, threads)}


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else a library prints (NCCL banners ...) went to stderr."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def main():
    global _REAL_STDOUT
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)                      # fd 1 -> stderr for the whole run (C libraries included)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
