#!/bin/bash
timeout 200 python -m pytest tests/test_gpu_api.py -m gpu -x -q -k "signal" > gpurun_out/r02_pytest_signal.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/r02_pytest_signal.log
TR="timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
$TR bench.py --gpus 2 --steps 60 --warmup 10 --workload c4 --scaling strong > gpurun_out/r02_bench_n2_c4s_flag.json 2> gpurun_out/r02_bench_n2_c4s_flag.err; echo rc=$?
$TR bench.py --gpus 2 --steps 60 --warmup 10 > gpurun_out/r02_bench_n2_c2_flag.json 2> gpurun_out/r02_bench_n2_c2_flag.err; echo rc=$?
python - <<'PY'
import json
for f in ("r02_bench_n2_c4s_flag", "r02_bench_n2_c2_flag"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f)); g = d.get("gather") or {}; gg = d.get("gathered") or {}
        print(f, "value %.0f ms %.4f kernel %.4f | pipelined %.0f = %.4f ms per frame | gathered %.0f %s | check %s" % (
            d["value"], d["ms_per_step"], d["kernel_ms_per_launch"], d["pipelined"]["value"], 1e3 / d["pipelined"]["frames_per_s"], gg.get("value") or 0, gg.get("check"), d["gather_check"]))
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/%s.err" % f).read()[-2000:])
PY
