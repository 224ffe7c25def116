#!/usr/bin/env python
"""Summarise `ncu --page source --csv` (one kernel) by warp role: control region (between
USETMAXREG.DEALLOC and TRY_ALLOC) vs softmax region, top stall reasons and hottest instructions.
usage: ncu_stalls.py SRC.csv [RAW.csv] [NTOP]"""
import csv
import sys


def load(fn):
    rows = list(csv.reader(open(fn)))
    hdr = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        if len(r) < len(hdr) or not r[idx['# Samples']].isdigit():
            break
        out.append(r)
    return hdr, idx, out


def main():
    hdr, idx, rows = load(sys.argv[1])
    if len(sys.argv) > 2:
        raw = list(csv.reader(open(sys.argv[2])))
        ri = {h: i for i, h in enumerate(raw[0])}
        for w in ['gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
                  'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
                  'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
                  'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
                  'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_active.avg',
                  'sm__cycles_elapsed.avg']:
            if w in ri:
                print(w, [r[ri[w]] for r in raw[2:]])
    stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    start = [i for i, r in enumerate(rows) if 'USETMAXREG.TRY_ALLOC' in r[idx['Source']]][0]
    dealloc = [i for i, r in enumerate(rows) if 'USETMAXREG.DEALLOC' in r[idx['Source']]][0]
    ntop = int(sys.argv[3]) if len(sys.argv) > 3 else 16
    for name, (a, b) in (('control', (dealloc, start)), ('softmax', (start, len(rows)))):
        sub = rows[a:b]
        n = sum(int(r[idx['# Samples']]) for r in sub)
        agg = {s: sum(int(r[idx[s]]) for r in sub) for s in stalls}
        nm = sum(int(r[idx['Instructions Executed']]) for r in sub if 'UTCHMMA' in r[idx['Source']])
        ni = sum(int(r[idx['Instructions Executed']]) for r in sub)
        print(name, 'samples', n, 'instr', ni, 'UTCHMMA', nm,
              [(k[6:], v) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:9]])
        for r in sorted(sub, key=lambda r: -int(r[idx['# Samples']]))[:ntop]:
            st = sorted(((s[6:], int(r[idx[s]])) for s in stalls), key=lambda x: -x[1])[:3]
            print('   ', r[idx['# Samples']].rjust(6), r[idx['Instructions Executed']].rjust(8),
                  r[idx['Source']].strip()[:64].ljust(64), st)


if __name__ == "__main__":
    main()
