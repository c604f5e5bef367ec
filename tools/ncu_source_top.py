"""Top instructions by stall samples of one kernel in an `ncu --page source --csv` export, with their main stall reasons.
python tools/ncu_source_top.py file.csv kernel_index [top=40]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = int(sys.argv[2]); top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
ks = []; cur = None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'hdr': None, 'rows': []}; ks.append(cur); continue
    if cur is None: continue
    if r and r[0] == 'Address': cur['hdr'] = r; continue
    if cur['hdr'] and len(r) >= len(cur['hdr']) - 2: cur['rows'].append(r)
k = ks[want]; h = k['hdr']; idx = {n: i for i, n in enumerate(h)}
stall = [n for n in h if n.startswith('stall_') and 'Not Issued' not in n]
S = idx['# Samples']
order = sorted(range(len(k['rows'])), key=lambda i: -int(k['rows'][i][S]))[:top]
for i in sorted(order):
    r = k['rows'][i]
    st = sorted(((int(r[idx[s]]), s[6:]) for s in stall), reverse=True)[:3]
    print('%5d %-60s smp %6d exec %9s  %s' % (i, r[1].strip()[:60], int(r[S]), r[idx['Instructions Executed']],
                                             ' '.join('%s=%d' % (n, c) for c, n in st if c)))
