"""Group an ncu `--metrics gpu__time_duration.sum --csv` log by kernel name: count, total us, share."""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1], newline='') as fh:
    lines = [l for l in fh if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get('Metric Name') == 'gpu__time_duration.sum':
        v = float(r['Metric Value'].replace(',', ''))
        unit = r.get('Metric Unit', 'ns')
        us = v / 1e3 if unit.startswith('n') else v if unit.startswith('u') else v * 1e3
        name = re.sub(r'\(.*', '', r['Kernel Name'])
        rows.append((name, us))
agg = defaultdict(lambda: [0, 0.0])
for n, us in rows:
    agg[n][0] += 1
    agg[n][1] += us
total = sum(v[1] for v in agg.values())
print('%d launches, %.1f us total' % (len(rows), total))
for n, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print('%6.1f%%  %9.1f us  %5d x  %s' % (100 * us / total, us, c, n[:110]))
