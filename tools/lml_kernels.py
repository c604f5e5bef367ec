"""Per-kernel totals of one batched LML+gradient evaluation from an ncu launch list
(ncu --metrics gpu__time_duration.sum --csv --log-file X python tools/fit_once.py n B)."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    a = agg.setdefault(r[4].split('(')[0], [0, 0.0])
    a[0] += 1
    a[1] += float(r[-1])
tot = sum(v[1] for v in agg.values())
for k, v in agg.items():
    print(k.ljust(34), '%3d launches %9.1f us %5.1f%%' % (v[0], v[1] / 1e3, 100 * v[1] / tot))
print('total %.3f ms' % (tot / 1e6))
