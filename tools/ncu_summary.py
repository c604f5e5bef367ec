"""Summarise an .ncu-rep (raw + source pages) into the few numbers the design notes quote."""
import csv
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'launch__block_size', 'launch__grid_size', 'sm__issue_active.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_elapsed.max',
        'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active']
for r in rows[2:]:
    print('== kernel', r[hdr.index('Kernel Name')][:60])
    for k in keep:
        if k in hdr:
            i = hdr.index(k)
            print('  %-70s %-14s %s' % (k, units[i], r[i]))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
sec = []
for r in rows[2:]:
    if len(r) < len(hdr) or r[0] in ('Kernel Name', 'Address'):
        break
    sec.append(r)
tot = sum(int(r[ix['# Samples']]) for r in sec)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
agg = Counter()
for r in sec:
    for s in stalls:
        agg[s] += int(r[ix[s]])
print('samples', tot)
for s, v in agg.most_common(8):
    print('  %-24s %5.1f%%' % (s, 100.0 * v / tot))
ops, ex = Counter(), Counter()
for r in sec:
    toks = [o for o in r[ix['Source']].split() if not o.startswith('@')]
    op = toks[0].split('.')[0] if toks else '?'
    ops[op] += int(r[ix['# Samples']])
    ex[op] += int(r[ix['Instructions Executed']])
print('samples by opcode')
for k, v in ops.most_common(12):
    print('  %-10s %5.1f%%  executed %d' % (k, 100.0 * v / tot, ex[k]))
if len(sys.argv) > 2:
    top = sorted(range(len(sec)), key=lambda i: -int(sec[i][ix['# Samples']]))[:int(sys.argv[2])]
    for i in sorted(top):
        r = sec[i]
        s = {h: int(r[ix[h]]) for h in stalls}
        main = sorted(s.items(), key=lambda t: -t[1])[:2]
        print(i, r[ix['Source']][:64].ljust(64), r[ix['# Samples']].rjust(7), '%.2f%%' % (100 * int(r[ix['# Samples']]) / tot), main)
