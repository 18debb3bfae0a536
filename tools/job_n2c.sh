#!/bin/bash
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
$TR bench.py --gpus 2 --steps 60 --warmup 10 > gpurun_out/r02_bench_n2_c2_c.json 2> gpurun_out/r02_bench_n2_c2_c.err; echo rc=$?
$TR bench.py --gpus 2 --steps 30 --warmup 5 --workload c4 --scaling strong > gpurun_out/r02_bench_n2_c4s_c.json 2> gpurun_out/r02_bench_n2_c4s_c.err; echo rc=$?
python - <<'PY'
import json
for f in ("r02_bench_n2_c2_c", "r02_bench_n2_c4s_c"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f)); g = d.get("gather") or {}
        print(f, "value %.0f ms %.4f pipelined %.0f | gathered %s | gather %s | e2e %.0f check %s" % (d["value"], d["ms_per_step"], d["pipelined"]["value"], json.dumps(d.get("gathered"))[:400], json.dumps(g)[:200], d["e2e"]["value"], d["gather_check"]))
        print(json.dumps(d["run"])[:600])
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/%s.err" % f).read()[-2500:])
PY
