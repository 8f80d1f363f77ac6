#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time per kernel name."""
import csv, collections, re, sys
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
lines = [l for l in open(path) if not l.startswith('==')]
tot = collections.Counter(); cnt = collections.Counter(); n = 0
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum': continue
    v = float(row['Metric Value'].replace(',', '')); u = row['Metric Unit']
    v = v / 1000 if u == 'ns' else (v * 1000 if u == 'ms' else v)
    name = re.sub(r'\(.*', '', row['Kernel Name'])[:90]
    tot[name] += v; cnt[name] += 1; n += 1
print(f"launches {n}  total {sum(tot.values()):.1f} us")
for k, v in tot.most_common(top):
    print(f"{v:10.1f} us {cnt[k]:5d} x {v/cnt[k]:8.1f}  {k}")
