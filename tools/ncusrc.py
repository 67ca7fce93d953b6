#!/usr/bin/env python
"""usage: tools/ncusrc.py <file.ncu-rep> [top]  -- executed warp instructions per CUDA source line (first kernel),
from ncu's correlated cuda,sass source page (needs -lineinfo and --import-source on)."""
import csv, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
per = collections.Counter(); text = {}; fn = ""; samples = collections.Counter()
for r in rows:
    if r and r[0] == "File Path": fn = r[1].split("/")[-1]; continue
    if not r or r[0] in ("Function Name", "Line No"): continue
    if r[0].isdigit() and len(r) > 7:
        key = (fn, int(r[0])); text[key] = r[1].strip()
        try: per[key] += int(r[7]); samples[key] += int(r[4])
        except ValueError: pass
tot = sum(per.values()); ts = sum(samples.values())
print("total warp instr", tot, "samples", ts)
byfile = collections.Counter()
for (f, l), v in per.items(): byfile[f] += v
for f, v in byfile.most_common(): print(f"  {f:24s} {v:10d} {100*v/tot:5.1f}%")
for k, v in per.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 60):
    print(f"{k[0]:18s}:{k[1]:5d} {v:9d} {100*v/tot:5.1f}%  smp {100*samples[k]/max(ts,1):4.1f}%  {text[k][:110]}")
