#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for x in csv.DictReader(lines):
    name = re.sub(r"alegpu::|<unnamed>::|\(anonymous namespace\)::", "", x["Kernel Name"])
    v = float(x["Metric Value"].replace(",", ""))
    u = x["Metric Unit"]
    v = v / 1000 if u in ("ns", "nsecond") else v * 1000 if u in ("ms", "msecond") else v
    a = agg.setdefault(name + " grid=" + x["Grid Size"], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"total {tot:.1f} us over {sum(a[0] for a in agg.values())} launches")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{a[1]:10.1f} us {a[0]:5d} x {a[1]/a[0]:8.2f} us {100*a[1]/tot:5.1f}%  {k[:170]}")
