"""Average duration per kernel name of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
kn, mv = h.index("Kernel Name"), h.index("Metric Value")
d = collections.defaultdict(list)
for r in rows[hi + 1:]:
    if len(r) > mv:
        try:
            d[r[kn].split("(")[0][:60]].append(float(r[mv].replace(",", "")))
        except ValueError:
            pass
for k, v in d.items():
    print(f"{k:60s} n={len(v):3d} avg={sum(v) / len(v) / 1e3:9.1f} us")
