#!/usr/bin/env python
"""Per-opcode executed-instruction counts per pixel from an ncu report's
source page:  python tools/ncu_ops.py report.ncu-rep <pixels> [kernel-index]"""
import collections
import csv
import subprocess
import sys

rep, px = sys.argv[1], float(sys.argv[2])
want = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv',
                      '--print-source', 'sass'], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(out.splitlines()))
blocks, cur = [], None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1], 'rows': []}
        blocks.append(cur)
    elif r and r[0] == 'Address':
        cur['hdr'] = r
    elif cur is not None and len(r) > 10:
        cur['rows'].append(r)
b = blocks[want]
h = b['hdr']
i_src, i_ex, i_s = h.index('Source'), h.index('Instructions Executed'), \
    h.index('# Samples')
ops, samp, tot, stot = collections.Counter(), collections.Counter(), 0, 0
for r in b['rows']:
    parts = r[i_src].split()
    op = parts[1] if parts[0].startswith('@') else parts[0]
    op = op.split('.')[0]
    n = int(r[i_ex])
    ops[op] += n
    tot += n
    samp[op] += int(r[i_s])
    stot += int(r[i_s])
print(b['name'])
print('warp instructions %d = %.2f per pixel' % (tot, tot * 32 / px))
for op, n in ops.most_common(28):
    print('%-10s %6.2f /px  %5.1f %% of samples' %
          (op, n * 32 / px, 100.0 * samp[op] / max(stot, 1)))
