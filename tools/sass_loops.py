#!/usr/bin/env python
"""Loop bodies of one kernel in a cuobjdump -sass dump: for every backward branch, the instruction count and opcode
histogram of [target, branch].  Used to read the per-window instruction budget of the demodulator's hot loop here,
before spending GPU time (B200_PROFILING.md: check the SASS first).
   cuobjdump -sass build/demod_cert.o | python tools/sass_loops.py demod_cert_kernelILi7ELb0 [min_instructions]"""
import re
import sys
from collections import Counter

name = sys.argv[1]
min_len = int(sys.argv[2]) if len(sys.argv) > 2 else 200
ins = []          # (addr, opcode, text)
on = False
for line in sys.stdin:
    if "Function :" in line:
        on = name in line
        continue
    if not on:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if not m:
        continue
    addr, text = int(m.group(1), 16), m.group(2).strip()
    t = text.split()
    op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    ins.append((addr, op.split(".")[0], text))
index = {a: i for i, (a, _, _) in enumerate(ins)}
print("kernel %s: %d instructions" % (name, len(ins)))
loops = []
for i, (a, op, text) in enumerate(ins):
    if op == "BRA":
        m = re.search(r"0x([0-9a-f]+)", text)
        if m:
            tgt = int(m.group(1), 16)
            if tgt <= a and tgt in index and i - index[tgt] + 1 >= min_len:
                loops.append((index[tgt], i))
for lo, hi in loops:
    body = ins[lo:hi + 1]
    c = Counter(op for _, op, _ in body)
    print("loop 0x%x..0x%x: %d instructions" % (body[0][0], body[-1][0], len(body)))
    print("   " + ", ".join("%s %d" % kv for kv in c.most_common(28)))
