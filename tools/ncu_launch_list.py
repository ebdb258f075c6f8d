#!/usr/bin/env python
"""ncu --metrics gpu__time_duration.sum --csv launch list -> per-kernel totals and shares.  usage: ncu_launch_list.py in.csv out.txt "header" """
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
rows = list(csv.reader(lines))
hdr = rows[0]
ik, iv, iu = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0}
tot, n = collections.OrderedDict(), collections.Counter()
for r in rows[1:]:
    k = re.sub(r'^void ', '', r[ik]).split('(')[0].replace('rt::', '').replace('<unnamed>::', '')
    tot[k] = tot.get(k, 0.0) + float(r[iv].replace(',', '')) * scale.get(r[iu], 1e-6); n[k] += 1
T = sum(tot.values())
with open(sys.argv[2], 'w') as f:
    f.write(sys.argv[3] + "\n")
    f.write(f"{'kernel':64s} {'launches':>8s} {'total_ms':>10s} {'avg_us':>9s} {'share':>7s}\n")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        f.write(f"{k:64s} {n[k]:8d} {v:10.3f} {v / n[k] * 1e3:9.1f} {v / T:7.3f}\n")
print(open(sys.argv[2]).read())
