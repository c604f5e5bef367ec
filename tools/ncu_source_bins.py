"""Stall samples of one kernel of an `ncu --page source --csv` export in bins of SASS rows.
python tools/ncu_source_bins.py file.csv kernel_index [bin=500]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
want = int(sys.argv[2]); B = int(sys.argv[3]) if len(sys.argv) > 3 else 500
ks = []; cur = None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'hdr': None, 'rows': []}; ks.append(cur); continue
    if cur is None: continue
    if r and r[0] == 'Address': cur['hdr'] = r; continue
    if cur['hdr'] and len(r) >= len(cur['hdr']) - 2: cur['rows'].append(r)
k = ks[want]; h = k['hdr']; idx = {n: i for i, n in enumerate(h)}
stall = [n for n in h if n.startswith('stall_') and 'Not Issued' not in n]
S = idx['# Samples']; E = idx['Instructions Executed']
for b0 in range(0, len(k['rows']), B):
    rs = k['rows'][b0:b0 + B]
    smp = sum(int(r[S]) for r in rs); ex = sum(int(r[E]) for r in rs)
    if smp < 200: continue
    c = collections.Counter()
    for r in rs:
        for s in stall: c[s[6:]] += int(r[idx[s]])
    fp = sum(int(r[E]) for r in rs if r[1].split() and [x for x in r[1].split() if not x.startswith('@')][0][:4] in ('DFMA', 'DADD', 'DMUL'))
    print('%5d-%5d smp %7d exec %11d fp64 %11d  %s' % (b0, b0 + B, smp, ex, fp, ' '.join('%s=%d' % (n, v) for n, v in c.most_common(5))))
