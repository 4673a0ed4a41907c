"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = None
agg = collections.defaultdict(list)
for r in rows:
    if len(r) > 5 and r[0] == "ID":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d.get("Metric Name") == "gpu__time_duration.sum":
            v = float(d["Metric Value"].replace(",", ""))
            u = d["Metric Unit"]
            if u in ("nsecond", "ns"):
                v /= 1000
            elif u in ("msecond", "ms"):
                v *= 1000
            agg[d["Kernel Name"][:70]].append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:70s} n={len(v):4d} avg={sum(v)/len(v):8.2f} us  total={sum(v)/1000:7.3f} ms  {100*sum(v)/tot:5.1f}%")
print(f"sum of launches: {tot/1000:.3f} ms (cold-cache, serialised under ncu)")
