"""Lists the backward-branch loops of a SASS dump with their instruction mix
(cuobjdump -sass -fun <kernel> file.o > x.sass; python tools/sass_loops.py x.sass)."""
import re
import sys
from collections import Counter

ins = []
for l in open(sys.argv[1]):
    m = re.search(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
minlen = int(sys.argv[2]) if len(sys.argv) > 2 else 20
for a, t in ins:
    m = re.search(r'BRA\S*\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)', t)
    if m and int(m.group(1), 16) < a:
        tgt = int(m.group(1), 16)
        body = [x for x in ins if tgt <= x[0] <= a]
        ops = []
        for _, s in body:
            p = s.split()
            op = p[1] if p[0].startswith('@') else p[0]
            ops.append(op.split('.')[0])
        c = Counter(ops)
        if len(body) >= minlen:
            print(hex(tgt), hex(a), len(body), dict(c.most_common(16)))
