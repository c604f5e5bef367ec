"""Stall-sample summary of an `ncu --page source --csv` export: per kernel, samples by stall reason and by opcode.
python tools/ncu_source_summary.py gpurun_out/pb_source.csv"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
ks = []; cur = None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1], 'hdr': None, 'rows': []}; ks.append(cur); continue
    if cur is None: continue
    if r and r[0] == 'Address': cur['hdr'] = r; continue
    if cur['hdr'] and len(r) >= len(cur['hdr']) - 2: cur['rows'].append(r)
for ki, k in enumerate(ks):
    h = k['hdr']; idx = {n: i for i, n in enumerate(h)}
    stall = [n for n in h if n.startswith('stall_') and 'Not Issued' not in n]
    S = idx['# Samples']
    total = sum(int(r[S]) for r in k['rows'])
    print('kernel', ki, 'instructions', len(k['rows']), 'samples', total)
    tot = collections.Counter()
    for r in k['rows']:
        for s in stall: tot[s] += int(r[idx[s]])
    print('  ', tot.most_common(10))
    op = collections.Counter(); opi = collections.Counter()
    for r in k['rows']:
        o = [x for x in r[1].split() if not x.startswith('@')]
        name = o[0].rstrip(';') if o else '?'
        op[name] += int(r[S]); opi[name] += int(r[idx['Instructions Executed']])
    for n, c in op.most_common(28): print('  %-24s samples %7d  warp-inst %11d' % (n, c, opi[n]))
