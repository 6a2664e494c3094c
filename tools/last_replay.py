#!/usr/bin/env python
"""Per-kernel durations of the LAST graph replay in an ncu launch list (gpu__time_duration.sum csv) of tools/infer_once.py"""
import csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
def us(r):
    v = float(r['Metric Value'].replace(',', '')); u = r['Metric Unit']
    return v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
names = [r['Kernel Name'] for r in rows]
fps = [i for i, n in enumerate(names) if 'fps_multi' in n]
start = fps[-1]                       # the sampling chain is the first node of the graph
tot = 0
for r in rows[start:]:
    print("%8.2f  %s" % (us(r), r['Kernel Name'][:110])); tot += us(r)
print("%8.2f  total (serialised, cold cache) over %d launches" % (tot, len(rows) - start))
