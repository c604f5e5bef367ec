"""Summarise an `ncu --page source --csv --print-source sass` export: per kernel, warp-stall samples by
opcode and by stall reason, and the instructions with most samples."""
import csv
import collections
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
kernels, cur, header = [], None, None
for row in csv.reader(open(path)):
    if not row:
        continue
    if row[0] == 'Kernel Name':
        cur = {'name': row[1], 'rows': []}
        kernels.append(cur)
        header = None
    elif row[0] == 'Address':
        header = row
    elif cur is not None and header is not None:
        cur['rows'].append(dict(zip(header, row)))
for k in kernels:
    rows = k['rows']
    total = sum(int(r['# Samples']) for r in rows)
    print('=== %s: %d instructions, %d samples' % (k['name'][:60], len(rows), total))
    by_op = collections.Counter()
    execd = collections.Counter()
    for r in rows:
        op = r['Source'].split()
        op = op[1] if op and op[0].startswith('@') else (op[0] if op else '?')
        op = op.split('.')[0]
        by_op[op] += int(r['# Samples'])
        execd[op] += int(r['Instructions Executed'])
    print('samples by opcode:', ', '.join('%s %.1f%%' % (o, 100.0 * c / total) for o, c in by_op.most_common(14)))
    tot_exec = sum(execd.values())
    print('executed by opcode:', ', '.join('%s %.1f%%' % (o, 100.0 * c / tot_exec) for o, c in execd.most_common(14)), '(total %d)' % tot_exec)
    stalls = collections.Counter()
    for r in rows:
        for key, val in r.items():
            if key.startswith('stall_') and not key.endswith('(Not Issued)') and val not in ('', '-'):
                stalls[key] += int(val)
    print('stall reasons:', ', '.join('%s %.1f%%' % (s, 100.0 * c / total) for s, c in stalls.most_common(10)))
    print('top instructions:')
    for r in sorted(rows, key=lambda r: -int(r['# Samples']))[:top]:
        why = sorted(((int(v), key) for key, v in r.items() if key.startswith('stall_') and not key.endswith('(Not Issued)') and v not in ('', '-', '0')), reverse=True)[:2]
        print('  %6d  %-70s %s' % (int(r['# Samples']), r['Source'].strip()[:70], ' '.join('%s=%d' % (b, a) for a, b in why)))
