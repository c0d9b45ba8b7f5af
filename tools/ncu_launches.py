#!/usr/bin/env python
"""Summarise an ncu launch list (--csv --log-file ...): per launch the kernel and every metric collected.
    python tools/ncu_launches.py gpurun_out/launches.csv [name-width]"""
import csv
import sys
from collections import OrderedDict

rows = list(csv.reader(open(sys.argv[1])))
width = int(sys.argv[2]) if len(sys.argv) > 2 else 44
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
ix = {n: h.index(n) for n in ("Kernel Name", "Metric Name", "Metric Value", "ID")}
d = OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) < len(h):
        continue
    d.setdefault((int(r[ix["ID"]]), r[ix["Kernel Name"]][:width]), {})[r[ix["Metric Name"]]] = r[ix["Metric Value"]]
for (i, name), v in d.items():
    print(i, name, " ".join(f"{k.split('__')[-1]}={val}" for k, val in v.items()))
