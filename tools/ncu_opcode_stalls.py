#!/usr/bin/env python3
"""Aggregate the `ncu --page source --csv` export of one kernel by SASS opcode: executed instructions, stall samples and the
dominant stall reasons. usage: ncu_opcode_stalls.py source.csv"""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; data = rows[hi + 1:]
ix = {h: i for i, h in enumerate(hdr)}
tot = defaultdict(float); n = defaultdict(float); st = defaultdict(lambda: defaultdict(float)); allsamp = 0
KEYS = ('stall_wait', 'stall_math', 'stall_not_selected', 'stall_selected', 'stall_short_sb', 'stall_long_sb', 'stall_dispatch',
        'stall_branch_resolving', 'stall_no_inst', 'stall_mio', 'stall_lg')
for r in data:
    if len(r) < len(hdr) or r[0] == "Address": continue
    op = r[ix['Source']].strip().split()
    if not op: continue
    o = op[1] if op[0].startswith('@') else op[0]
    o = o.split('.')[0]
    try:
        ex = float(r[ix['Instructions Executed']] or 0); s = float(r[ix['# Samples']] or 0)
    except ValueError:
        continue
    tot[o] += s; n[o] += ex; allsamp += s
    for k in KEYS:
        st[o][k] += float(r[ix[k]] or 0)
print('total samples %d, total warp instructions %.4e' % (allsamp, sum(n.values())))
fp = sum(n[k] for k in ('DFMA', 'DMUL', 'DADD', 'DSETP', 'DMNMX'))
print('FP64 warp instructions %.4e (%.1f%%)' % (fp, 100 * fp / sum(n.values())))
for o, s in sorted(tot.items(), key=lambda kv: -kv[1])[:18]:
    print('%-8s inst=%.3e samples=%6d (%4.1f%%) ' % (o, n[o], s, 100 * s / allsamp),
          {k.replace('stall_', ''): int(v) for k, v in st[o].items() if v > 0.03 * s})
