#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
mkdir -p gpurun_out/steps_rgba gpurun_out/steps_rgb
BENCH_DUMP_STEPS=gpurun_out/steps_rgba $TR bench.py --gpus 8 --steps 200 --warmup 10 --gather-format rgba8 > gpurun_out/r02_bench_n8_c2_rgba_b.json 2> gpurun_out/r02_bench_n8_c2_rgba_b.err; echo rc=$?
BENCH_DUMP_STEPS=gpurun_out/steps_rgb $TR bench.py --gpus 8 --steps 200 --warmup 10 > gpurun_out/r02_bench_n8_c2_b.json 2> gpurun_out/r02_bench_n8_c2_b.err; echo rc=$?
python - <<'PY'
import numpy as np
for d in ("steps_rgba", "steps_rgb"):
    a = np.stack([np.load("gpurun_out/%s/steps_rank%d.npy" % (d, r)) for r in range(8)])   # ranks x steps
    print(d, "per-rank mean us:", np.round(a.mean(1) * 1e3, 1), "median:", np.round(np.median(a, 1) * 1e3, 1), "p90:", np.round(np.percentile(a, 90, 1) * 1e3, 1), "max:", np.round(a.max(1) * 1e3, 1))
    worst = a.max(0); idx = np.argsort(worst)[-8:]
    print("   slowest steps (index: per-rank us):", "; ".join("%d: %s" % (i, np.round(a[:, i] * 1e3, 0).astype(int).tolist()) for i in sorted(idx)))
    print("   pose of step i = i % 3 (rank r offset r): mean by i%3 for rank 1:", [round(float(a[1, k::3].mean() * 1e3), 1) for k in range(3)], "rank 0:", [round(float(a[0, k::3].mean() * 1e3), 1) for k in range(3)])
PY
