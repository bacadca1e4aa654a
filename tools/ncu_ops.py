#!/usr/bin/env python
"""Dynamic opcode mix of the profiled kernel: python tools/ncu_ops.py rep [divisor]"""
import csv, subprocess, sys, io, collections, re
rep = sys.argv[1]; div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
hdr = None; ops = collections.Counter(); tot = 0
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "Address": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].startswith("0x"):
        inst = int(r[hdr.index("Instructions Executed")])
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[1].strip())
        ops[m.group(2) if m else r[1][:10]] += inst; tot += inst
print("warp-inst", tot)
for k, v in ops.most_common(40): print(f"{k:10s} {100*v/tot:5.1f}%  {v/div:10.3g}")
