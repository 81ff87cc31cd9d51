#!/usr/bin/env python
"""Per CUDA source line instruction counts and stall samples (needs -lineinfo).
usage: tools/ncu_lines.py file.ncu-rep [top]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur_file = "?"; hdr = None; agg = collections.Counter(); samp = collections.Counter(); text = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; iI = hdr.index("Instructions Executed"); iS = hdr.index("# Samples"); continue
    if hdr is None or len(r) < len(hdr) or r[0] == "": continue
    k = (cur_file, int(r[0])); text[k] = r[1].strip()
    try: agg[k] += int(r[iI]); samp[k] += int(r[iS])
    except ValueError: pass
tot = sum(agg.values()); ts = sum(samp.values())
print("total warp instructions", tot, "samples", ts)
for k, n in agg.most_common(top):
    print(f"{100.0*n/tot:5.1f}% inst {100.0*samp[k]/max(ts,1):5.1f}% smp  {k[0]}:{k[1]:<4d} {text.get(k,'')[:90]}")
