"""Builder tool: aggregate an `ncu --csv --metrics gpu__time_duration.sum` launch list by kernel name."""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1], newline='') as f:
    lines = [l for l in f if not l.startswith('==')]
rd = csv.DictReader(lines)
tot = defaultdict(lambda: [0, 0.0])
for r in rd:
    if r.get('Metric Name') != 'gpu__time_duration.sum':
        continue
    name = re.sub(r'\(.*', '', r['Kernel Name'])
    v = float(r['Metric Value'].replace(',', ''))
    unit = r.get('Metric Unit', 'ns')
    if unit in ('us', 'usecond'):
        v *= 1e3
    elif unit in ('ms', 'msecond'):
        v *= 1e6
    tot[name][0] += 1
    tot[name][1] += v
total = sum(v[1] for v in tot.values())
print(f'{"kernel":60s} {"launches":>8s} {"total ms":>10s} {"avg us":>9s} {"share":>6s}')
for k, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f'{k[:60]:60s} {c:8d} {t / 1e6:10.3f} {t / c / 1e3:9.2f} {100 * t / total:5.1f}%')
print(f'{"total":60s} {sum(v[0] for v in tot.values()):8d} {total / 1e6:10.3f}')
