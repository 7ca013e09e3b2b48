#!/usr/bin/env python3
"""Sum `ncu --page source` per-line counts per device function of scan_kernel.cu.
usage: ncu_phases.py src.csv scan_kernel.cu"""
import csv, re, sys
src = open(sys.argv[2]).read().split('\n')
funcs = []   # (start_line, name)
for i, l in enumerate(src, 1):
    m = re.match(r'^(?:__device__|__global__|kvq_)[^(]*?(\w+)\(', l) or re.match(r'^(kvq_\w+)\(', l)
    if l.startswith('__device__') or l.startswith('__global__') or l.startswith('kvq_scan_kernel') or l.startswith('kvq_long_kernel'):
        name = re.findall(r'(\w+)\(', l)
        if name: funcs.append((i, name[0]))
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; out = []
for r in rows:
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr and r and r[0].isdigit():
        d = dict(zip(hdr, r)); num = lambda k: int(d[k]) if d[k].lstrip('-').isdigit() else 0
        out.append((int(r[0]), num('Instructions Executed'), num('Thread Instructions Executed'), num('# Samples')))
tot = sum(o[1] for o in out); ts = sum(o[3] for o in out)
acc = {}
for ln, ie, te, sm in out:
    name = 'header'
    for s0, n in funcs:
        if s0 <= ln: name = n
    a = acc.setdefault(name, [0, 0, 0]); a[0] += ie; a[1] += te; a[2] += sm
print('total warp instr %d, samples %d' % (tot, ts))
for n, (i, t, s) in sorted(acc.items(), key=lambda kv: -kv[1][0]):
    print('%-22s %11d %5.1f%% lanes %5.1f stall %5.1f%%' % (n, i, 100. * i / tot, t / max(1, i), 100. * s / max(1, ts)))
