#!/usr/bin/env python3
"""Turns `ncu --set full` captures of the timed render kernel into profiles/ncu_summary.json, the file bench.py reads its executed
instruction counts and DRAM traffic from (nothing is typed into bench.py).

  tools/ncu_summary.py --commit <sha> c2=gpurun_out/r02_c2.ncu-rep c4=gpurun_out/r02_c4.ncu-rep ...

Per workload: per-launch MEANS over the captured launches (tools/prof.sh captures one launch per cycled pose), plus the per-launch
rows, and a text rendering under profiles/<tag>_ncu_<workload>.txt.  Runs here (no GPU): `ncu -i` only reads the report."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PIXELS = {"c1": 400 * 400, "c2": 1920 * 1080, "c3": 1920 * 1080, "c4": 3840 * 2160, "c5": 7680 * 4320, "c5s": 3840 * 2160}
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__sass_average_branch_targets_threads_uniform.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active"]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    args = sys.argv[1:]
    commit, tag = "unknown", "r02"
    while args and args[0].startswith("--"):
        if args[0] == "--commit": commit = args[1]
        elif args[0] == "--tag": tag = args[1]
        args = args[2:]
    out_path = os.path.join(ROOT, "profiles", "ncu_summary.json")
    summary = json.load(open(out_path)) if os.path.exists(out_path) else {}
    for spec in args:
        wl, rep = spec.split("=", 1)
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, launches = rows[0], rows[1], rows[2:]
        col = {h: i for i, h in enumerate(hdr)}
        kname = launches[0][col["Kernel Name"]]

        def metric(name, row):
            v = num(row[col[name]])
            return None if v is None else v * SCALE.get(units[col[name]], 1.0)

        per = []
        for r in launches:
            d = {k: metric(k, r) for k in KEEP if k in col}
            d["thread_inst"] = d["smsp__inst_executed.sum"] * d["smsp__thread_inst_executed_per_inst_executed.ratio"]
            per.append(d)
        mean = lambda k: sum(p[k] for p in per) / len(per)
        clk = mean("sm__cycles_elapsed.avg.per_second")        # GHz during the capture
        summary[wl] = {
            "file": "profiles/%s_ncu_%s.txt" % (tag, wl), "kernel": kname, "commit": commit, "launches": len(per),
            "duration_us_per_launch": mean("gpu__time_duration.sum"),
            "dram_bytes_per_launch": mean("dram__bytes_read.sum") + mean("dram__bytes_write.sum"),
            "warp_inst_per_launch": mean("smsp__inst_executed.sum"), "thread_inst_per_launch": mean("thread_inst"),
            "thread_inst_per_pixel": mean("thread_inst") / PIXELS.get(wl, 1),
            "lanes_per_inst": mean("smsp__thread_inst_executed_per_inst_executed.ratio"),
            "issue_active_pct": mean("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "active_over_elapsed": mean("smsp__cycles_active.avg") / mean("sm__cycles_elapsed.avg"),
            "sm_ghz_during_capture": clk,
            # executed thread-instructions / capture duration / (148 SM x 128 lanes x capture clock): the capture's own utilisation
            "frac_actual_in_capture": sum(p["thread_inst"] / (p["gpu__time_duration.sum"] * 1e-6) / (148 * 128 * p["sm__cycles_elapsed.avg.per_second"] * 1e9) for p in per) / len(per),
            "registers": mean("launch__registers_per_thread"),
        }
        with open(os.path.join(ROOT, "profiles", "%s_ncu_%s.txt" % (tag, wl)), "w") as f:
            f.write("# ncu --set full --clock-control none, %d launches of the TIMED kernel (one per cycled pose), commit %s\n# %s\n# source: %s\n" %
                    (len(per), commit, kname, os.path.basename(rep)))
            for k in KEEP + ["thread_inst"]:
                if k in per[0] and per[0][k] is not None:
                    f.write("%-95s %s\n" % (k + (" [%s]" % units[col[k]] if k in col and units[col[k]] else ""), " | ".join("%.6g" % p[k] for p in per)))
            f.write("\nsummary: " + json.dumps(summary[wl], indent=1) + "\n")
        print(wl, json.dumps(summary[wl]))
    json.dump(summary, open(out_path, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
