#!/bin/bash
TR="timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
$TR bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/r02_bench_n8_c2.json 2> gpurun_out/r02_bench_n8_c2.err; echo rc=$?
$TR bench.py --gpus 8 --steps 120 --warmup 10 --workload c4 --scaling strong > gpurun_out/r02_bench_n8_c4s.json 2> gpurun_out/r02_bench_n8_c4s.err; echo rc=$?
$TR bench.py --gpus 8 --steps 8 --warmup 3 --workload c5 > gpurun_out/r02_bench_n8_c5.json 2> gpurun_out/r02_bench_n8_c5.err; echo rc=$?
python - <<'PY'
import json
for f in ("r02_bench_n8_c2", "r02_bench_n8_c4s", "r02_bench_n8_c5"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f)); g = d.get("gather") or {}; gg = d.get("gathered") or {}
        print(f, "value %.0f ms %.4f median %.4f kernel %.4f | pipelined %.0f = %.4f ms/frame | gather %.0f of %.0f GB/s | gathered %.0f (%s) | e2e %.0f check %s" % (
            d["value"], d["ms_per_step"], d["ms_per_step_median"], d["kernel_ms_per_launch"], d["pipelined"]["value"], 1e3 / d["pipelined"]["frames_per_s"], g.get("achieved_gbs") or 0, g.get("ingest_peak_gbs") or 0,
            gg.get("value") or 0, gg.get("check"), d["e2e"]["value"], d["gather_check"]))
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/%s.err" % f).read()[-1500:])
PY
