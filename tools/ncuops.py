#!/usr/bin/env python
"""usage: tools/ncuops.py <file.ncu-rep> [top]  -- executed warp instructions by opcode (first kernel)"""
import csv, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
ex = collections.Counter(); n = 0; hdr = None
for r in rows:
    if r and r[0] == "Kernel Name":
        n += 1
        if n > 1: break
        continue
    if r and r[0] == "Address": hdr = r; ie = hdr.index("Instructions Executed"); continue
    if hdr and len(r) > ie:
        t = r[1].split()
        if not t: continue
        op = (t[1] if t[0].startswith('@') else t[0]).rstrip(';')
        ex[op] += int(r[ie])
tot = sum(ex.values())
print("total warp instr", tot)
for k, v in ex.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 40):
    print(f"{k:28s} {v:10d} {100*v/tot:5.1f}%")
