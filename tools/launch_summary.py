"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import csv
import json
import re
import sys
from collections import OrderedDict

path = sys.argv[1]
rows = []
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.reader(lines)
hdr = next(rd)
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = OrderedDict()
for r in rd:
    name = r[ik]
    m = re.search(r"phased_kernel<afe::(\w+(<\w+>)?)", name)
    short = m.group(1) if m else re.sub(r"\(.*", "", name)[:60]
    v = float(r[iv].replace(",", ""))
    v = v / 1e3 if r[iu] in ("ns", "nsecond") else v          # -> us
    a = agg.setdefault(short, [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
out = {"source": path, "total_us": round(tot, 1), "kernels": {k: {"launches": a[0], "us": round(a[1], 1),
       "share": round(a[1] / tot, 4)} for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])}}
print(json.dumps(out, indent=1))
