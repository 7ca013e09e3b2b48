#!/usr/bin/env python3
"""Summarise `ncu --page source --csv --print-source cuda,sass` per CUDA source line:
warp instructions executed, share, lane efficiency, stall samples.  usage: ncu_lines.py file.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None; out = []
for r in rows:
    if r and r[0] == 'Line No':
        hdr = r; continue
    if hdr and r and r[0].isdigit():
        d = dict(zip(hdr, r))
        num = lambda k: int(d[k]) if d[k].lstrip('-').isdigit() else 0
        ie = num('Instructions Executed'); te = num('Thread Instructions Executed')
        out.append((int(r[0]), ie, te, num('# Samples'), r[1].strip()))
tot = sum(o[1] for o in out); ts = sum(o[3] for o in out)
print('total warp instr %d, samples %d' % (tot, ts))
print('%5s %12s %6s %6s %7s  %s' % ('line', 'warp_instr', 'share', 'lanes', 'stall%', 'source'))
for o in sorted(out, key=lambda o: -o[1])[:top]:
    print('%5d %12d %5.1f%% %6.1f %6.1f%%  %s' % (o[0], o[1], 100. * o[1] / tot, o[2] / max(1, o[1]), 100. * o[3] / max(1, ts), o[4][:90]))
