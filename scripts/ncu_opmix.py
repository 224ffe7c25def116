#!/usr/bin/env python
"""Opcode mix (executed warp-instructions) of the softmax region of an ncu source-page CSV.
usage: ncu_opmix.py SRC.csv [per_step_divisor]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[2:]:
    if len(r) < len(hdr) or not r[idx['# Samples']].isdigit():
        break
    body.append(r)
start = [i for i, r in enumerate(body) if 'USETMAXREG.TRY_ALLOC' in r[idx['Source']]][0]
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
mix, samp = collections.Counter(), collections.Counter()
for r in body[start:]:
    toks = r[idx['Source']].strip().split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    parts = op.split('.')
    op = parts[0] + ('.' + parts[1] if parts[0] in ('MUFU', 'LDTM', 'STS', 'SYNCS', 'F2FP') and len(parts) > 1 else '')
    mix[op] += int(r[idx['Instructions Executed']])
    samp[op] += int(r[idx['# Samples']])
tot = sum(mix.values())
print('total', tot / div)
for op, n in mix.most_common(24):
    print(f'{op:24s} {n / div:10.1f} {100 * n / tot:5.1f}%  samples {samp[op]}')
