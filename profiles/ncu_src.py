#!/usr/bin/env python3
"""Per-CUDA-source-line totals from `ncu --page source --csv --print-source sass,cuda` (first launch of each file view).
usage: ncu_src.py file.csv [topN]"""
import csv, sys, os
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
seen = set(); cur = None; take = False; hdr = None
out = []
for r in rows:
    if not r: continue
    if r[0] == 'File Path':
        cur = os.path.basename(r[1]); take = cur not in seen; seen.add(cur); continue
    if r[0] == 'Line No': hdr = {h: i for i, h in enumerate(r)}; continue
    if r[0] == 'Function Name' or not take or hdr is None: continue
    if r[0].strip().isdigit():
        try:
            ie = int(r[7]); te = int(r[8]); ss = int(r[6])
        except ValueError:
            continue
        out.append((ie, te, ss, cur, int(r[0]), r[1].strip()[:100]))
ti = sum(o[0] for o in out); tt = sum(o[1] for o in out); ts = sum(o[2] for o in out)
print("total warp-inst %d thread-inst %d (avg lanes %.1f) samples %d" % (ti, tt, tt / max(1, ti), ts))
byfile = {}
for o in out: byfile[o[3]] = byfile.get(o[3], 0) + o[0]
print({k: "%.1f%%" % (100.0 * v / ti) for k, v in byfile.items()})
for o in sorted(out, key=lambda x: -x[0])[:top]:
    print("%5.1f%% inst %5.1f%% smp lanes %4.1f | %s:%d %s" % (100.0 * o[0] / ti, 100.0 * o[2] / max(1, ts), o[1] / max(1, o[0]), o[3], o[4], o[5]))
