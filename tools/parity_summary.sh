#!/bin/bash
# one line per config: mismatches vs cap, rms, actual steps   (args passed to parity_report.py)
python "$(dirname "$0")/parity_report.py" "$@" 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(d['cfg'],d['pose'],d['mismatch_px'],'cap',d['cap_px'],d['rms255'],d['actual_steps'])
"
