#!/usr/bin/env python
"""Per-kernel breakdown of ONE forward taken from an ncu launch list of bench.py (forward #4 = the timed one)."""
import csv, re, sys
from collections import OrderedDict
path = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lines = [l for l in open(path) if not l.startswith('==')]
rows = [r for r in csv.DictReader(lines) if r['Metric Name'] == 'gpu__time_duration.sum']
names = [re.sub(r'\(.*', '', r['Kernel Name']) for r in rows]
names = [re.sub(r'^void ', '', n).replace('ttsb::', '')[:60] for n in names]
starts = [i for i, n in enumerate(names) if 'embed_pe_ln' in n]
s, e = starts[which], starts[which + 1]
agg = OrderedDict(); tot = 0
for r, n in zip(rows[s:e], names[s:e]):
    v = float(r['Metric Value'].replace(',', '')); a = agg.setdefault((n, r['Grid Size']), [0, 0.0]); a[0] += 1; a[1] += v; tot += v
print(e - s, 'launches', round(tot / 1e3, 1), 'us')
for (k, g), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{t/1e3:9.1f} us {100*t/tot:5.1f}% n={n:3d} avg {t/n/1e3:7.1f}  {k} {g}")
