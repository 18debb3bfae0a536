#!/bin/bash
TR="timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519"
$TR bench.py --gpus 4 --steps 100 --warmup 10 > gpurun_out/r02_bench_n4_c2.json 2> gpurun_out/r02_bench_n4_c2.err; echo rc=$?
$TR bench.py --gpus 4 --steps 20 --warmup 3 --impl reference > gpurun_out/r02_bench_n4_ref.json 2> gpurun_out/r02_bench_n4_ref.err; echo ref rc=$?
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/r02_bench_n4_c2.json")); gg = d.get("gathered") or {}
    print("n4 c2 value %.0f ms %.4f kernel %.4f pipelined %.0f | gathered %.0f (%s) | e2e %.0f" % (d["value"], d["ms_per_step"], d["kernel_ms_per_launch"], d["pipelined"]["value"], gg.get("value") or 0, gg.get("check"), d["e2e"]["value"]))
except Exception as e:
    print("FAILED", e); print(open("gpurun_out/r02_bench_n4_c2.err").read()[-1500:])
try:
    d = json.load(open("gpurun_out/r02_bench_n4_ref.json")); print("ref under torchrun n4:", d["value"], d["n_gpus"])
except Exception as e:
    print("ref FAILED", e); print(open("gpurun_out/r02_bench_n4_ref.err").read()[-800:])
PY
