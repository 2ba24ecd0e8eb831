"""Reads an ncu launch list (--metrics gpu__time_duration.sum --csv) and prints the 0-based ordinal, among the launches whose
kernel name matches argv[2], of the longest one (used as `-s` of the follow-up `ncu --set full -k regex:... -c 1` capture)."""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], errors="replace")))
hdr = None
best, best_i, i = -1.0, 0, 0
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    d = dict(zip(hdr, r))
    if d.get("Metric Name") != "gpu__time_duration.sum" or not re.search(sys.argv[2], d["Kernel Name"]):
        continue
    v = float(d["Metric Value"].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(d["Metric Unit"], 1.0)
    if v > best:
        best, best_i = v, i
    i += 1
print(best_i)
