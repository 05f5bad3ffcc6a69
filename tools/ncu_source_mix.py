#!/usr/bin/env python
"""Dynamic instruction mix + stall samples per opcode from `ncu -i X.ncu-rep --page source --csv` (first kernel in the file)."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
isrc, iex, ist = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Warp Stall Sampling (All Samples)')
data = []
for r in rows[2:]:
    if len(r) <= max(isrc, iex, ist) or not r[iex].isdigit():
        if data: break
        continue
    data.append(r)
tot = sum(int(r[iex]) for r in data); tots = sum(int(r[ist]) for r in data)
print('kernel', rows[0][1][:70]); print('total warp-inst', tot, 'stall samples', tots, 'sass lines', len(data))
h, hs = collections.Counter(), collections.Counter()
for r in data:
    t = r[isrc].split()
    op = (t[1] if t[0].startswith('@') else t[0]).rstrip(';')
    op = '.'.join(op.split('.')[:2]) if op.startswith(('LDS', 'STG', 'LDG', 'STS', 'BAR', 'SYNCS')) else op.split('.')[0]
    h[op] += int(r[iex]); hs[op] += int(r[ist])
for op, c in h.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f"{op:14s} {c / tot * 100:5.1f}% inst  {hs[op] / tots * 100:5.1f}% stall samples")
if len(sys.argv) > 3:  # top stalled lines
    for r in sorted(data, key=lambda r: -int(r[ist]))[:int(sys.argv[3])]:
        print(r[ist].rjust(6), r[iex].rjust(9), r[isrc][:100])
