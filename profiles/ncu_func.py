#!/usr/bin/env python3
"""Attributes warp instructions of `ncu --page source --csv --print-source sass,cuda` to source functions
(line ranges found by scanning the .cuh/.cu files for 'VRT_HD'/'static VRT_HD'/'__global__' definitions).
usage: ncu_func.py src.csv csrc_dir"""
import csv, os, re, sys, bisect, collections
rows = list(csv.reader(open(sys.argv[1]))); csrc = sys.argv[2]
def funcs_of(path):
    out = []
    for i, l in enumerate(open(path), 1):
        m = re.match(r'\s*(?:static\s+)?(?:VRT_HD|VRT_D|inline|__global__|__device__ __forceinline__)\b.*?(\w+)\s*\(', l)
        if m and not l.strip().startswith('//'): out.append((i, m.group(1)))
    return out
cache = {}
cur = None; take = False; seen = set(); agg = collections.defaultdict(lambda: [0, 0, 0])
for r in rows:
    if not r: continue
    if r[0] == 'File Path':
        cur = os.path.basename(r[1]); take = cur not in seen; seen.add(cur); continue
    if r[0] in ('Line No', 'Function Name') or not take: continue
    if r[0].strip().isdigit():
        try: ie, te, ss = int(r[7]), int(r[8]), int(r[6])
        except ValueError: continue
        if cur not in cache:
            p = os.path.join(csrc, cur); cache[cur] = funcs_of(p) if os.path.exists(p) else []
        fl = cache[cur]; i = bisect.bisect_right([x[0] for x in fl], int(r[0])) - 1
        k = cur[:-4][:12] + ':' + (fl[i][1] if i >= 0 else '?')
        agg[k][0] += ie; agg[k][1] += te; agg[k][2] += ss
tot = sum(v[0] for v in agg.values()); ts = sum(v[2] for v in agg.values())
print("warp-inst %.1fM (line-attributed; inlined lines may be counted in several views)" % (tot / 1e6))
for k, v in sorted(agg.items(), key=lambda x: -x[1][0])[:32]:
    print("%5.1f%% inst %5.1f%% smp %6.1fM lanes %4.1f %s" % (100 * v[0] / tot, 100 * v[2] / max(ts, 1), v[0] / 1e6, v[1] / max(1, v[0]), k))
