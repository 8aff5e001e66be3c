#!/usr/bin/env python
"""Top stalled SASS instructions per kernel from `ncu -i X --page source --csv` output (stdin or file)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 15
sections, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        sections.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
for s in sections:
    hdr, data = s["hdr"], s["data"]
    i_src, i_s = hdr.index("Source"), hdr.index("# Samples")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(d[i_s]) for d in data)
    print("====", s["name"][:140], "samples", tot, "instrs", len(data))
    agg = {}
    for d in data:
        for st in stalls:
            agg[st[6:]] = agg.get(st[6:], 0) + int(d[hdr.index(st)])
    print("   stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    for n, d in sorted(enumerate(data), key=lambda x: -int(x[1][i_s]))[:topn]:
        st = {x[6:]: int(d[hdr.index(x)]) for x in stalls if int(d[hdr.index(x)]) > 0}
        print(f"   {n:5d} {d[i_s]:>6} {d[i_src][:64].strip():64s} {st}")
