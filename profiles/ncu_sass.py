#!/usr/bin/env python3
"""Lists SASS instructions (address order, deduplicated) with execution counts above a threshold.
usage: ncu_sass.py src.csv min_count"""
import csv, os, sys
rows = list(csv.reader(open(sys.argv[1]))); thr = int(sys.argv[2])
cur = None; take = False; seen = set(); line = None; out = {}
for r in rows:
    if not r: continue
    if r[0] == 'File Path':
        cur = os.path.basename(r[1]); take = cur not in seen; seen.add(cur); continue
    if r[0] in ('Line No', 'Function Name') or not take: continue
    if r[0].strip().isdigit(): line = (cur, int(r[0])); continue
    if r[2].startswith('0x'):
        try: a = int(r[2], 16); out.setdefault(a, (line, r[3].strip(), int(r[7]), int(r[8]), int(r[6])))
        except ValueError: pass
ks = sorted(out); base = ks[0]
print("# %d SASS instructions, %d warp-inst executed" % (len(ks), sum(out[k][2] for k in ks)))
for k in ks:
    l, s, ie, te, ss = out[k]
    if ie >= thr: print("%5x %-20s %9d %4.1f %4d %s" % (k - base, "%s:%d" % (l[0][:-4][:13], l[1]), ie, te / max(ie, 1), ss, s[:90]))
