#!/usr/bin/env python
"""Executed warp instructions and stall samples of one kernel split at backward branches: for every loop found in the
SASS of the report (innermost first), its share of the kernel's executed instructions, executions of its first
instruction and instructions per iteration.
usage: tools/ncu_regions.py file.ncu-rep [min_loop_instructions]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]; min_len = int(sys.argv[2]) if len(sys.argv) > 2 else 150
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; ins = []
for r in rows:
    if r and r[0] == "Address": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    ins.append((int(r[0], 16), r[1].strip(), int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("# Samples")])))
base = ins[0][0]
idx = {a: i for i, (a, _, _, _) in enumerate(ins)}
total = sum(x[2] for x in ins); tsamp = sum(x[3] for x in ins)
print("kernel: %d SASS instructions, %d executed warp instructions, %d samples" % (len(ins), total, tsamp))
loops = []
for i, (a, text, n, s) in enumerate(ins):
    if re.search(r"\bBRA\b", text):
        m = re.search(r"0x([0-9a-f]+)", text)
        if m:
            t = int(m.group(1), 16)
            if t not in idx: t += base
            if t in idx and idx[t] <= i and i - idx[t] + 1 >= min_len: loops.append((idx[t], i))
for lo, hi in sorted(loops, key=lambda p: p[1] - p[0]):
    ex = sum(x[2] for x in ins[lo:hi + 1]); sm = sum(x[3] for x in ins[lo:hi + 1])
    it = max(1, ins[lo][2])
    print("loop +0x%x..+0x%x (%d instr): %.1f%% of executed, %.1f%% of samples, %d entries of the first instruction, %.0f executed per entry"
          % (ins[lo][0] - base, ins[hi][0] - base, hi - lo + 1, 100.0 * ex / total, 100.0 * sm / max(1, tsamp), it, ex / it))
