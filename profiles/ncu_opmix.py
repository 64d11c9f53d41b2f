#!/usr/bin/env python
"""Instruction mix + stall samples per SASS opcode from `ncu -i X.ncu-rep --page source --csv --kernel-id :::N > f.csv`."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if 'Instructions Executed' in r)
iS, iE, iSm = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
ops, samp, st = collections.Counter(), collections.Counter(), collections.Counter()
tot = 0
seen = set()
for r in rows:
    if len(r) < len(hdr) or not r[iE].isdigit() or not r[0].startswith('0x') or r[0] in seen:
        continue
    seen.add(r[0])
    parts = r[iS].strip().split()
    op = parts[1] if parts[0].startswith('@') else parts[0]
    op = op.split('.')[0]
    n = int(r[iE]); ops[op] += n; tot += n; samp[op] += int(r[iSm])
    for s in stalls:
        st[s] += int(r[hdr.index(s)])
print('total warp instructions', tot, ' SASS lines', len(seen))
for op, n in ops.most_common(30):
    print(f'{op:10s} {n:10d} {100*n/tot:5.1f}%  samples {samp[op]}')
ts = sum(st.values())
print('stall samples:', [(k, round(100*v/ts, 1)) for k, v in st.most_common(10)])
