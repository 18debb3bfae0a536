#!/bin/bash
# final-ish single-GPU lines for every workload + C5 ncu
for w in c2 c3 c4; do python bench.py --workload $w --steps 100 --warmup 10 > gpurun_out/r02_bench_n1_$w.json 2> gpurun_out/r02_bench_n1_$w.err; echo $w rc=$?; done
python bench.py --workload c5 --steps 16 --warmup 3 > gpurun_out/r02_bench_n1_c5.json 2> gpurun_out/r02_bench_n1_c5.err; echo c5 rc=$?
python bench.py --workload c1 --steps 100 --warmup 10 > gpurun_out/r02_bench_n1_c1.json 2> gpurun_out/r02_bench_n1_c1.err; echo c1 rc=$?
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_ref_c2.json 2> gpurun_out/r02_bench_ref_c2.err; echo ref rc=$?
python - <<'PY'
import json
for w in ("c2", "c3", "c4", "c5", "c1"):
    try:
        d = json.load(open("gpurun_out/r02_bench_n1_%s.json" % w)); r = d["roofline"]
        print(w, "value %.0f ms %.4f median %.4f min %.4f | frac %s eff %.2f | e2e %.0f | cpu %s | clocks %s" % (d["value"], d["ms_per_step"], d["ms_per_step_median"], d["ms_per_step_min"], r.get("frac"), r["effective"]["frac"], d["e2e"]["value"], (d.get("cpu_baseline") or {}).get("value"), json.dumps(d["clocks"])[:160]))
    except Exception as e:
        print(w, "FAILED", e); print(open("gpurun_out/r02_bench_n1_%s.err" % w).read()[-1500:])
d = json.load(open("gpurun_out/r02_bench_ref_c2.json")); print("ref", d["value"], d["ms_per_step"], d["cpu_baseline"]["sample"])
PY
