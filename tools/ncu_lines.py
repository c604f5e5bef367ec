"""Per-source-line warp-stall samples from `ncu --page source --csv --print-source cuda,sass`."""
import csv
import collections
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
cur_file = cur_fn = hdr = None
agg = collections.defaultdict(collections.Counter)
src = {}
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
        continue
    if r[0] == 'Function Name':
        cur_fn = r[1][:24]
        continue
    if r[0] == 'Line No':
        hdr = r
        continue
    if hdr is None or not r[0].isdigit():
        continue
    d = dict(zip(hdr[2:], r[2:]))
    key = (cur_fn, cur_file, int(r[0]))
    src.setdefault(key, r[1])
    if r[2] != '-':        # per-instruction rows repeat what the line row sums up
        continue
    try:
        agg[key]['samples'] += int(d['# Samples'])
        agg[key]['exec'] += int(d['Instructions Executed'])
    except (KeyError, ValueError):
        pass
for fn in sorted(set(k[0] for k in agg)):
    tot = sum(v['samples'] for k, v in agg.items() if k[0] == fn)
    print('=====', fn, tot)
    items = [(v['samples'], k) for k, v in agg.items() if k[0] == fn and v['samples'] > 0]
    for s, k in sorted(items, reverse=True)[:top]:
        print('%6.2f%%  %s:%d  %s' % (100.0 * s / tot, k[1], k[2], src[k].strip()[:100]))
