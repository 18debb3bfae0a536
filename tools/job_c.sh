#!/bin/bash
# GPU job: full GPU test suite, A/B timing of in-tree vs build/* variants, sun-field build cost
python -m pytest tests -m gpu -x -q --durations=5 -s > gpurun_out/r02_pytest_c.log 2>&1; echo pytest rc=$?; grep -E "mismatching px|passed|failed|Error|^E " gpurun_out/r02_pytest_c.log | tail -12
python tools/ab_time.py --cases c2,c3,c4,c4b - $(ls build/*/libvrt.so 2>/dev/null) > gpurun_out/r02_ab_c.txt 2>&1; cat gpurun_out/r02_ab_c.txt
python - <<'PY' > gpurun_out/r02_sun_field_cost.txt 2>&1
import sys, time
sys.path.insert(0, "tests")
import numpy as np, common, voxelraytracer_b200 as vrt
ctx = vrt.Context(0)
for sc, W, H in (("terrain128", 1920, 1080), ("refraction128", 1920, 1080), ("terrain128_glass", 3840, 2160)):
    grid, n, atlas, ts = common.scene(sc); ctx.upload_grid(grid, n); ctx.upload_atlas(atlas, ts)
    pose = {"terrain128": "terrain_inside", "refraction128": "refraction_close", "terrain128_glass": "glass_wall"}[sc]
    out = ctx.device_alloc(W * H * 4)
    def run(tods, reps=40):
        t = []
        for i in range(reps + 5):
            p = common.make_params(W, H, sc, pose, time_of_day=tods[i % len(tods)])
            ctx.synchronize(); t0 = time.perf_counter(); ctx.render_to(p, out); ctx.synchronize(); dt = time.perf_counter() - t0
            if i >= 5: t.append(dt * 1e3)
        return float(np.median(t))
    static = run([0.0]); moving = run([0.01 * k for k in range(50)])
    ctx.set_option(vrt.OPT_SUN_FIELD, 0); off = run([0.0]); ctx.set_option(vrt.OPT_SUN_FIELD, 1)
    print("%s %dx%d  wall ms per frame (render + sync): sun field, static sun %.4f | sun moving every frame (field rebuilt) %.4f | sun field off %.4f" % (sc, W, H, static, moving, off))
PY
cat gpurun_out/r02_sun_field_cost.txt
