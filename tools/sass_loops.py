#!/usr/bin/env python
"""List the loops (backward branches) of a kernel in a cuobjdump -sass dump with their opcode mix.
usage: cuobjdump -sass -fun <mangled> lib.so | python tools/sass_loops.py [min_instructions]"""
import re, sys
from collections import Counter
minlen = int(sys.argv[1]) if len(sys.argv) > 1 else 40
ins = []
for line in sys.stdin:
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
addr_index = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t)
    if m:
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in addr_index:
            j = addr_index[tgt]
            body = ins[j:i + 1]
            if len(body) < minlen:
                continue
            c = Counter()
            for _, tt in body:
                tt = re.sub(r"^@!?U?P\d+\s+", "", tt)
                c[tt.split()[0].split(".")[0]] += 1
            print(f"loop 0x{tgt:x}..0x{a:x}: {len(body)} instructions")
            print("   ", ", ".join(f"{k} {v}" for k, v in c.most_common(14)))
