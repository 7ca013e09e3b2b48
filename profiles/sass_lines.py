#!/usr/bin/env python3
"""Static SASS census of one kernel: instructions per source line / per function, by opcode.
usage: sass_lines.py <file.sass from `nvdisasm --print-line-info -c x.cubin`> <kernel substring> [lo hi]"""
import re, sys
from collections import Counter
path, pat = sys.argv[1], sys.argv[2]
lo, hi = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (0, 10**9)
lines = open(path).read().split('\n')
start = next(i for i, l in enumerate(lines) if l.startswith('.text.') and pat in l)
cur = None; cnt = Counter(); ops = {}
n = 0
for l in lines[start + 1:]:
    if l.startswith('\t.section') or l.startswith('.text.'):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m2 = re.match(r'\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)', l)
    if m2 and cur:
        cnt[cur] += 1; ops.setdefault(cur, Counter())[m2.group(2).split('.')[0]] += 1; n += 1
print('static instructions', n)
for k in sorted(cnt):
    if lo <= k[1] <= hi:
        print('%-22s %5d  %4d  %s' % (k[0][:22], k[1], cnt[k], dict(ops[k])))
