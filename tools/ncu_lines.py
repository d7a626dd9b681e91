#!/usr/bin/env python
"""Warp-stall samples and executed instructions per CUDA source line, from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > X.csv` (needs -lineinfo + --import-source on)."""
import csv, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
cur = fn = hdr = None
agg = {}
for r in csv.reader(open(path)):
    if not r: continue
    if r[0] == 'File Path': cur = r[1]; continue
    if r[0] == 'Function Name': fn = r[1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr and cur and len(r) == len(hdr) and r[0].isdigit():
        si = hdr.index('# Samples'); ie = hdr.index('Instructions Executed')
        try: s = int(r[si] or 0); ins = int(r[ie] or 0)
        except ValueError: continue
        a = agg.setdefault((fn[:70], cur.split('/')[-1], int(r[0])), [0, 0, r[1][:110]]); a[0] += s; a[1] += ins
for f in sorted(set(k[0] for k in agg)):
    tot = sum(v[0] for k, v in agg.items() if k[0] == f); toti = sum(v[1] for k, v in agg.items() if k[0] == f)
    print('==', f, 'samples', tot, 'warp-instructions', toti)
    for k, v in sorted(((k, v) for k, v in agg.items() if k[0] == f), key=lambda kv: -kv[1][0])[:topn]:
        print(f"  {k[1]:16s}:{k[2]:4d} samp {100*v[0]/max(tot,1):5.1f}% inst {100*v[1]/max(toti,1):5.1f}%  {v[2]}")
