#!/usr/bin/env python3
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
agg = collections.OrderedDict()
for r in csv.DictReader(lines):
    if r.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', r['Kernel Name'])[:90]
    v = float(r['Metric Value'].replace(',', ''))
    u = r['Metric Unit']
    v = v / 1000 if u == 'ns' else (v * 1000 if u == 'ms' else v)
    agg.setdefault(name, []).append(v)
tot = sum(sum(v) for v in agg.values())
print(f"total {tot/1000:.3f} ms over {sum(len(v) for v in agg.values())} launches")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:90s} n={len(v):4d} avg={sum(v)/len(v):9.1f}us tot={sum(v)/1000:8.3f}ms {100*sum(v)/tot:5.1f}%")
