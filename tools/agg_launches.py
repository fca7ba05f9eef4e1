"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name.

    python tools/agg_launches.py launches.csv            # the last complete training step (between the last two
                                                         # adam_kernel launches); falls back to everything
"""
import collections
import csv
import re
import sys

path = sys.argv[1]
lines = [l for l in open(path) if not l.startswith('==')]
rows = list(csv.DictReader(lines))
adam = [i for i, r in enumerate(rows) if 'adam_kernel' in r['Kernel Name']]
if len(adam) >= 2:
    rows = rows[adam[-2] + 1:adam[-1] + 1]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in rows:
    name = re.sub(r'\(.*', '', row['Kernel Name'])
    name = re.sub(r'^void ', '', name)
    v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v = v / 1000 if u == 'ns' else (v * 1000 if u == 'ms' else v)
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print('launches %d  total %.1f us (one step)' % (len(rows), tot))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    print('%-64s n=%4d  total %9.1f us  avg %7.2f us  %5.1f%%' % (k[:64], v[0], v[1], v[1] / v[0], 100 * v[1] / tot))
