#!/bin/bash
# 8-GPU job: weak (C2) and strong (C4) bench lines, both gather formats
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
$TR bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/r02_bench_n8_c2.json 2> gpurun_out/r02_bench_n8_c2.err; echo rc=$?
$TR bench.py --gpus 8 --steps 100 --warmup 10 --gather-format rgba8 > gpurun_out/r02_bench_n8_c2_rgba.json 2> gpurun_out/r02_bench_n8_c2_rgba.err; echo rc=$?
$TR bench.py --gpus 8 --steps 60 --warmup 10 --workload c4 --scaling strong > gpurun_out/r02_bench_n8_c4s.json 2> gpurun_out/r02_bench_n8_c4s.err; echo rc=$?
$TR bench.py --gpus 8 --steps 60 --warmup 10 --workload c4 --scaling strong --gather-format rgba8 > gpurun_out/r02_bench_n8_c4s_rgba.json 2> gpurun_out/r02_bench_n8_c4s_rgba.err; echo rc=$?
python bench.py --gpus 1 --steps 60 --warmup 10 --workload c4 --no-cpu-baseline > gpurun_out/r02_bench_n1_c4.json 2> gpurun_out/r02_bench_n1_c4.err; echo rc=$?
python - <<'PY'
import json
for f in ("r02_bench_n8_c2", "r02_bench_n8_c2_rgba", "r02_bench_n8_c4s", "r02_bench_n8_c4s_rgba", "r02_bench_n1_c4"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f))
        g = d.get("gather") or {}
        print(f, "value %.0f ms %.4f median %.4f min %.4f pipelined %.0f | gather %.0f of %.0f GB/s (pipelined %.0f) | e2e %.0f check %s" % (d["value"], d["ms_per_step"], d["ms_per_step_median"], d["ms_per_step_min"], d["pipelined"]["value"], g.get("achieved_gbs") or 0, g.get("ingest_peak_gbs") or 0, g.get("pipelined_gbs") or 0, d["e2e"]["value"], d["gather_check"]))
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/%s.err" % f).read()[-1500:])
PY
