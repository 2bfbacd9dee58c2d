#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per source line:
warp instructions executed, mean active lanes, stall samples.  usage: ncu_lines.py dump.csv [top]"""
import csv, sys, collections, os
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
fname = "?"; H = None
agg = collections.OrderedDict()
for r in rows:
    if not r: continue
    if r[0] in ("File Name", "File Path"): fname = os.path.basename(r[1]) or "?"; continue
    if r[0] == "Line No": H = r; continue
    if H is None or len(r) < len(H): continue
    try: ie = float(r[H.index("Instructions Executed")] or 0); te = float(r[H.index("Thread Instructions Executed")] or 0); sm = float(r[H.index("# Samples")] or 0)
    except ValueError: continue
    if not r[0]:  # a SASS row: already summed into its source-line row
        continue
    key = (fname, r[0]); a = agg.setdefault(key, [0, 0, 0, r[1].strip()[:90]])
    a[0] += ie; a[1] += te; a[2] += sm
tot = sum(a[0] for a in agg.values()) or 1; tots = sum(a[2] for a in agg.values()) or 1
print(f"total warp-inst {tot:.3e}, samples {tots:.0f}")
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), a in agg.items():
    byfile[f][0] += a[0]; byfile[f][1] += a[1]; byfile[f][2] += a[2]
for f, a in sorted(byfile.items(), key=lambda kv: -kv[1][0]): print(f"{f:28s} inst {a[0]/tot*100:5.1f}%  lanes {a[1]/max(a[0],1):5.1f}  samples {a[2]/tots*100:5.1f}%")
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
    print(f"{f}:{l:>4s} inst {a[0]/tot*100:5.1f}% lanes {a[1]/max(a[0],1):5.1f} samp {a[2]/tots*100:5.1f}% | {a[3]}")
