#!/usr/bin/env python
"""Top SASS instructions by warp-stall samples from `ncu --page source --csv` output."""
import csv, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
# several kernels may be concatenated: split on "Kernel Name" rows
blocks = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}; blocks.append(cur)
    elif cur is not None and r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] and len(r) == len(cur["hdr"]):
        cur["rows"].append(r)
for b in blocks:
    h = b["hdr"]; si = h.index("# Samples"); src = h.index("Source")
    stall = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[si] or 0) for r in b["rows"])
    print("==", b["name"][:90], "total samples", tot)
    agg = {}
    for r in b["rows"]:
        for i in stall:
            agg[h[i]] = agg.get(h[i], 0) + int(r[i] or 0)
    print("  stall totals:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    idx = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][si] or 0))[:topn]
    for i in sorted(idx):
        r = b["rows"][i]
        st = sorted(((int(r[k] or 0), h[k][6:]) for k in stall), reverse=True)[:2]
        print(f"  {i:5d} {int(r[si]):6d} {100*int(r[si])/max(tot,1):5.1f}%  {r[src][:90]:90s} {st}")
