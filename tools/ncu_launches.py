#!/usr/bin/env python
"""Prints an ncu --csv long-format log as one line per launch."""
import csv, sys
from collections import OrderedDict
lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
rows = list(csv.reader(lines))
hdr = rows[0]
ik, im, iv, iid, iu = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('ID'), hdr.index('Metric Unit')
d = OrderedDict()
for r in rows[1:]:
    d.setdefault((r[iid], r[ik].replace('void ', '').split('(')[0][:44]), OrderedDict())[r[im] + '[' + r[iu] + ']'] = r[iv]
for k, v in d.items():
    print(k[0].rjust(3), k[1].ljust(44), ' '.join(f"{m}={x}" for m, x in v.items()))
