#!/bin/bash
# Final single-GPU job at HEAD: GPU tests, smoke, A/B timing, bench lines for every workload, ncu captures + launch list of the timed kernels
timeout 600 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r02_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -9 gpurun_out/r02_pytest_final.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke_final.log
timeout 200 python tools/ab_time.py --cases c2,c3,c4,c4b - > gpurun_out/r02_ab_final.txt 2>&1; cat gpurun_out/r02_ab_final.txt
timeout 400 python bench.py --steps 100 --warmup 10 > gpurun_out/r02_bench_n1_c2.json 2> gpurun_out/r02_bench_n1_c2.err; echo c2 rc=$?
for w in c3 c4 c1; do timeout 300 python bench.py --workload $w --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r02_bench_n1_$w.json 2> gpurun_out/r02_bench_n1_$w.err; echo $w rc=$?; done
timeout 500 python bench.py --workload c5 --steps 16 --warmup 3 > gpurun_out/r02_bench_n1_c5.json 2> gpurun_out/r02_bench_n1_c5.err; echo c5 rc=$?
VRT_SUN_FIELD=2 timeout 400 python bench.py --workload c5 --steps 16 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_n1_c5_sunfield.json 2> gpurun_out/r02_bench_n1_c5_sunfield.err; echo c5sun rc=$?
timeout 400 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_ref_c2.json 2> gpurun_out/r02_bench_ref_c2.err; echo ref rc=$?
for w in c2 c3 c4; do bash tools/prof.sh r02_f_$w --workload $w; done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_f_launches_c2.csv python bench.py --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/r02_f_launches_c2.log 2>&1; echo launches rc=$?
python - <<'PY'
import json
for w in ("c2", "c3", "c4", "c1", "c5", "c5_sunfield"):
    try:
        d = json.load(open("gpurun_out/r02_bench_n1_%s.json" % w)); r = d["roofline"]
        print(w, "value %.0f ms %.4f median %.4f min %.4f kernel %.4f | frac %s eff %.2f | e2e %.0f | pipelined %.0f | cpu %s" % (d["value"], d["ms_per_step"], d["ms_per_step_median"], d["ms_per_step_min"], d["kernel_ms_per_launch"], r.get("frac"), r["effective"]["frac"], d["e2e"]["value"], d["pipelined"]["value"], json.dumps((d.get("cpu_baseline") or {}).get("value"))))
    except Exception as e:
        print(w, "FAILED", e); print(open("gpurun_out/r02_bench_n1_%s.err" % w).read()[-1200:])
try:
    d = json.load(open("gpurun_out/r02_bench_ref_c2.json")); print("ref", d["value"], d["ms_per_step"], d["cpu_baseline"]["sample"])
except Exception as e:
    print("ref FAILED", e)
PY
