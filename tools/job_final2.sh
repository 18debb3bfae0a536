#!/bin/bash
# Final single-GPU job at HEAD: GPU tests, smoke, bench lines for every workload (small files only)
timeout 600 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke.log
timeout 400 python bench.py --steps 100 --warmup 10 > gpurun_out/r02_bench_n1_c2.json 2> gpurun_out/r02_bench_n1_c2.err; echo c2 rc=$?
for w in c3 c4 c1; do timeout 300 python bench.py --workload $w --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r02_bench_n1_$w.json 2> gpurun_out/r02_bench_n1_$w.err; echo $w rc=$?; done
timeout 500 python bench.py --workload c5 --steps 16 --warmup 3 > gpurun_out/r02_bench_n1_c5.json 2> gpurun_out/r02_bench_n1_c5.err; echo c5 rc=$?
VRT_SUN_FIELD=2 timeout 400 python bench.py --workload c5 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_n1_c5_sunfield.json 2> gpurun_out/r02_bench_n1_c5_sunfield.err; echo c5sun rc=$?
timeout 400 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_ref_c2.json 2> gpurun_out/r02_bench_ref_c2.err; echo ref rc=$?
python - <<'PY'
import json
for w in ("c2", "c3", "c4", "c1", "c5", "c5_sunfield"):
    try:
        d = json.load(open("gpurun_out/r02_bench_n1_%s.json" % w)); r = d["roofline"]
        print(w, "value %.0f ms %.4f median %.4f min %.4f kernel %.4f | frac %s eff %.2f | e2e %.0f | pipelined %.0f | cpu %s | %s" % (d["value"], d["ms_per_step"], d["ms_per_step_median"], d["ms_per_step_min"], d["kernel_ms_per_launch"], r.get("frac"), r["effective"]["frac"], d["e2e"]["value"], d["pipelined"]["value"], json.dumps((d.get("cpu_baseline") or {}).get("value")), (r.get("ncu") or {}).get("kernel")))
    except Exception as e:
        print(w, "FAILED", e); print(open("gpurun_out/r02_bench_n1_%s.err" % w).read()[-1200:])
try:
    d = json.load(open("gpurun_out/r02_bench_ref_c2.json")); print("ref", d["value"], d["ms_per_step"], d["cpu_baseline"]["sample"])
except Exception as e:
    print("ref FAILED", e)
PY
