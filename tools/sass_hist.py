#!/usr/bin/env python
"""Opcode histogram of an address range of a cuobjdump -sass listing (static instruction mix per section).
usage: sass_hist.py listing.sass LO HI [LO HI ...]   (hex byte offsets)"""
import collections
import re
import sys

rx = re.compile(r"/\*([0-9a-f]{4,5})\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)")
ins = []
for line in open(sys.argv[1]):
    m = rx.search(line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
args = sys.argv[2:]
for lo, hi in zip(args[::2], args[1::2]):
    lo, hi = int(lo, 16), int(hi, 16)
    c = collections.Counter(op for a, op in ins if lo <= a < hi)
    print(f"[{lo:#x},{hi:#x}) {sum(c.values())} instr:", " ".join(f"{k}:{v}" for k, v in c.most_common()))
