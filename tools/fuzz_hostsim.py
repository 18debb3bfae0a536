#!/usr/bin/env python3
"""Randomised parity fuzz of the fast traversal (device headers compiled for the host) against the oracle: random camera
poses inside / outside / on the boundary of the chunk, random length caps, bounce limits, sun angles, frame shapes.
A quarter of the cases set the noise uniforms (careful instantiation: pinned ray construction + glass replay).
Every case runs in a child process with a timeout, so a walk that never terminates is reported, not waited for.
Bars per case: hit / bounce mismatches <= max(1, 0.01 % of the frame) + 2, and colour RMS <= 1/255 per channel.  The fuzz frames are
tiny (6-16 k pixels), where ONE grazing pixel that the hit cap allows can flip sky <-> stone and exceed the RMS bar by itself
(255 / sqrt(12800) = 2.3); the colour bar is therefore taken over the pixels whose hit / bounce entries agree - the grazing
pixels are what the hit cap counts - and the whole-frame RMS is reported beside it (`rms_full`).
  python tools/fuzz_hostsim.py [--cases N] [--seed S] [--timeout SEC]
TEST TOOLING (uses oracle/ and tests/hostsim)."""
import argparse, json, multiprocessing as mp, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

SCENES = ["glass_cube32", "terrain32", "refraction32", "terrain128", "terrain128_glass", "refraction128"]


def make_case(rng):
    sc = SCENES[int(rng.integers(len(SCENES)))]
    n = 32 if sc.endswith("32") else 128
    kind = int(rng.integers(5))
    h = n / 2.0
    if kind == 0:   pos = rng.uniform(-h, h, 3)                                # inside
    elif kind == 1: pos = rng.uniform(-3 * h, 3 * h, 3)                          # anywhere around
    elif kind == 2: pos = np.round(rng.uniform(-h, h, 3))                        # on integer planes (ties)
    elif kind == 3:                                                              # exactly on a chunk face
        pos = rng.uniform(-h, h, 3); pos[int(rng.integers(3))] = h * (1 if rng.random() < 0.5 else -1)
    else:           pos = rng.uniform(-h, h, 3) + 0.5 * np.sign(rng.uniform(-1, 1, 3)) * 1e-4
    rot = (float(rng.uniform(-1.55, 1.55)), float(rng.uniform(-np.pi, np.pi)), 0.0)
    if rng.random() < 0.15: rot = (float(rng.integers(-1, 2)) * np.pi / 2 * 0.999, float(rng.integers(-2, 3)) * np.pi / 2, 0.0)   # axis-aligned views
    W, H = [(160, 90), (97, 61), (256, 64), (64, 200)][int(rng.integers(4))]
    return dict(scene=sc, kind=["inside", "around", "lattice", "on_face", "near_lattice"][kind], pos=[float(x) for x in pos], rot=rot, W=W, H=H,
                max_len=float([100.0, 100.0, 7.5, 33.0, 400.0, 0.5][int(rng.integers(6))]),
                refl=int(rng.integers(0, 3)), trans=int(rng.integers(0, 4)), tod=float(rng.uniform(0, 50)),
                flat=bool(rng.random() < 0.3),
                noise=[float(x) for x in (rng.uniform(0, 0.1, 3) * (rng.random(3) < 0.6))] if rng.random() < 0.25 else [0.0, 0.0, 0.0],
                time=float(rng.integers(1, 500)))


def run_case(c, q):
    import common, voxelraytracer_b200 as vrt
    grid, n, atlas, ts = common.scene(c["scene"])
    p = common.make_params(c["W"], c["H"], c["scene"], (c["pos"], c["rot"]), vrt.FLAG_COLOR_ONLY if c["flat"] else 0, time_of_day=c["tod"], max_len=c["max_len"],
                           noise=tuple(c.get("noise", (0.0, 0.0, 0.0))), time=c.get("time", 1.0))
    p.maxReflections, p.maxTransparencies = c["refl"], c["trans"]
    o = common.oracle_render(p, grid, atlas, nthreads=2)
    g = common.hostsim_render(p, grid, atlas, 1, nthreads=2)
    e = common.hostsim_render(p, grid, atlas, 0, nthreads=2)
    differ = (g.hit != o.hit) | (g.bounces != o.bounces)
    bad = int(differ.sum())
    same = ~differ
    exact_ok = bool(np.array_equal(e.hit, o.hit) and np.array_equal(e.bounces, o.bounces) and np.array_equal(e.f32, o.f32))
    q.put(dict(bad=bad, exact_ok=exact_ok, rms=[float(x) for x in common.rms255(g.rgba8[same], o.rgba8[same])],
               rms_full=[float(x) for x in common.rms255(g.rgba8, o.rgba8)], npix=c["W"] * c["H"]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=200); ap.add_argument("--seed", type=int, default=1); ap.add_argument("--timeout", type=float, default=120.0)
    ap.add_argument("--jobs", type=int, default=max(1, (os.cpu_count() or 2) // 2))
    a = ap.parse_args()
    import common
    common.build_oracle(); common.build_hostsim()          # once, before the children (which would race on the same .so) start
    rng = np.random.default_rng(a.seed)
    cases = [make_case(rng) for _ in range(a.cases)]
    ctx = mp.get_context("fork")
    worst = dict(frac=0.0); worst_rms_full = 0.0; hung = []; inexact = []; over = []; lattice = []; total_bad = 0; total_px = 0
    pending = list(enumerate(cases)); running = []
    import time
    while pending or running:
        while pending and len(running) < a.jobs:
            i, c = pending.pop(0); q = ctx.Queue(); pr = ctx.Process(target=run_case, args=(c, q)); pr.start(); running.append((i, c, q, pr, time.time()))
        time.sleep(0.05)
        for item in list(running):
            i, c, q, pr, t0 = item
            if not q.empty():
                r = q.get(); pr.join(); running.remove(item)
                total_bad += r["bad"]; total_px += r["npix"]
                frac = r["bad"] / r["npix"]
                if not r["exact_ok"]: inexact.append((i, c))
                worst_rms_full = max(worst_rms_full, max(r["rms_full"]))
                if frac > worst["frac"]: worst = dict(frac=frac, case=i, bad=r["bad"], rms=r["rms"], cfg=c)
                if r["bad"] > max(1, int(1e-4 * r["npix"])) + 2 or max(r["rms"]) > 1.0: over.append((i, r["bad"], r["rms"], c))
            elif not pr.is_alive():
                pr.join(); running.remove(item); hung.append((i, "crashed rc=%s" % pr.exitcode, c))
            elif time.time() - t0 > a.timeout:
                pr.kill(); pr.join(); running.remove(item); hung.append((i, "timeout", c))
    print(json.dumps(dict(cases=a.cases, seed=a.seed, pixels=total_px, mismatch_px=total_bad, mismatch_frac=total_bad / max(total_px, 1),
                          hung_or_crashed=len(hung), exact_not_bit_exact=len(inexact), over_cap_cases=len(over), lattice_camera_over_cap=len(lattice),
                          noisy_cases=sum(1 for c in cases if any(c["noise"])), worst_rms_full=worst_rms_full, worst=worst)))
    for h in hung: print("HUNG/CRASHED", json.dumps(h))
    for h in inexact: print("EXACT MISMATCH", json.dumps(h))
    for h in over[:20]: print("OVER CAP", json.dumps(h))
    for h in lattice: print("LATTICE OVER CAP", json.dumps(h))
    return 1 if hung or inexact or over or lattice else 0


if __name__ == "__main__":
    sys.exit(main())
