"""Summarises an `ncu --csv` launch list of the wavefront kernels (per kernel: mean of each metric)."""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
acc = defaultdict(lambda: defaultdict(list))
for r in rows[1:]:
    k = r[ix["Kernel Name"]].split("(")[0].split("::")[-1]
    acc[k][r[ix["Metric Name"]]].append(float(r[ix["Metric Value"]].replace(",", "")))
for k, m in acc.items():
    print(k, {a: round(sum(v) / len(v), 2) for a, v in m.items()}, "launches", len(next(iter(m.values()))))
