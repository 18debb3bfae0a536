#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
$TR bench.py --gpus 2 --steps 60 --warmup 10 --workload c4 --scaling strong --gather dma > gpurun_out/r02_bench_n2_c4s_dma.json 2> gpurun_out/r02_bench_n2_c4s_dma.err; echo rc=$?
$TR bench.py --gpus 2 --steps 60 --warmup 10 --gather dma > gpurun_out/r02_bench_n2_c2_dma.json 2> gpurun_out/r02_bench_n2_c2_dma.err; echo rc=$?
$TR bench.py --gpus 2 --steps 60 --warmup 10 --gather none > gpurun_out/r02_bench_n2_c2_none.json 2> gpurun_out/r02_bench_n2_c2_none.err; echo rc=$?
python - <<'PY'
import json
for f in ("r02_bench_n2_c4s_dma", "r02_bench_n2_c2_dma", "r02_bench_n2_c2_none"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f)); g = d.get("gather") or {}
        print(f, "value %.0f ms %.4f pipelined %.0f | gather %.0f of %.0f GB/s (pipelined %.0f) | e2e %.0f check %s" % (d["value"], d["ms_per_step"], d["pipelined"]["value"], g.get("achieved_gbs") or 0, g.get("ingest_peak_gbs") or 0, g.get("pipelined_gbs") or 0, d["e2e"]["value"], d["gather_check"]))
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/%s.err" % f).read()[-1500:])
PY
