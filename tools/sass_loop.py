#!/usr/bin/env python
"""Instruction histogram of the innermost MUFU-bearing loop of each kernel in a
cuobjdump -sass listing (development tool)."""
import collections
import re
import sys

txt = open(sys.argv[1]).read()
pat = sys.argv[2] if len(sys.argv) > 2 else "."
for f in re.split(r'\n\s+Function : ', txt)[1:]:
    name = f.split('\n')[0]
    if not re.search(pat, name):
        continue
    ins = []
    for l in f.split('\n'):
        m = re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);', l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    loops = []
    for a, t in ins:
        mm = re.search(r'BRA\S*\s+(?:\S+,\s+)?0x([0-9a-f]+)', t)
        if mm and int(mm.group(1), 16) < a:
            loops.append((int(mm.group(1), 16), a))
    best = None
    for lo, hi in loops:
        if any('MUFU' in t for a, t in ins if lo <= a <= hi):
            if best is None or hi - lo < best[1] - best[0]:
                best = (lo, hi)
    if not best:
        continue
    body = [t for a, t in ins if best[0] <= a <= best[1]]
    c = collections.Counter(re.sub(r'^@!?U?P\d\s+', '', t).split()[0] for t in body)
    print(name[:110])
    print('  loop instrs', len(body), dict(c.most_common()))
