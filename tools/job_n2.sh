#!/bin/bash
# 2-GPU job: the 2-GPU IPC test, then weak (C2) and strong (C4) bench lines at N=2
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k two_gpus > gpurun_out/r02_pytest_n2.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/r02_pytest_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
$TR bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/r02_bench_n2_c2.json 2> gpurun_out/r02_bench_n2_c2.err; echo rc=$?
$TR bench.py --gpus 2 --steps 100 --warmup 10 --gather-format rgba8 > gpurun_out/r02_bench_n2_c2_rgba.json 2> gpurun_out/r02_bench_n2_c2_rgba.err; echo rc=$?
$TR bench.py --gpus 2 --steps 60 --warmup 10 --workload c4 --scaling strong > gpurun_out/r02_bench_n2_c4s.json 2> gpurun_out/r02_bench_n2_c4s.err; echo rc=$?
python - <<'PY'
import json
for f in ("r02_bench_n2_c2", "r02_bench_n2_c2_rgba", "r02_bench_n2_c4s"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f))
        print(f, "value %.0f ms %.4f median %.4f pipelined %.0f gather %s e2e %.0f check %s" % (d["value"], d["ms_per_step"], d["ms_per_step_median"], d["pipelined"]["value"], json.dumps(d.get("gather"))[:300], d["e2e"]["value"], d["gather_check"]))
    except Exception as e:
        print(f, "FAILED", e); print(open("gpurun_out/%s.err" % f).read()[-1500:])
PY
