#!/usr/bin/env python
"""Instruction share per source line / per SASS loop of an ncu report: python tools/ncu_regions.py rep [top]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, agg = None, {}
for r in csv.reader(io.StringIO(src)):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) > 8 and r[0].isdigit():
        try: inst, th, smp = int(r[7]), int(r[8]), int(r[6])
        except ValueError: continue
        agg[(cur, int(r[0]), r[1].strip()[:75])] = (inst, th, smp)
tot = sum(v[0] for v in agg.values()); tots = sum(v[2] for v in agg.values())
print("warp-inst", tot, "thread-inst", sum(v[1] for v in agg.values()), "samples", tots)
byfile = collections.Counter()
for k, v in agg.items(): byfile[k[0]] += v[0]
print({k: round(100 * v / tot, 1) for k, v in byfile.items()})
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100*v[0]/tot:5.1f}%i {100*v[2]/tots:5.1f}%s thr/inst={v[1]/max(v[0],1):5.1f} inst={v[0]:.3g} {k[0]}:{k[1]} {k[2]}")
