#!/usr/bin/env python
"""Summarise an ncu report per CUDA source line: python tools/ncu_lines.py rep.ncu-rep [top]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
want = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"]
want += [h for h in rows[0] if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")]
for h, u, v in zip(rows[0], rows[1], rows[2]):
    if h in want:
        print(f"{h:90s} {v} {u}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur, hdr, agg = None, None, []
for r in csv.reader(io.StringIO(src)):
    if len(r) >= 2 and r[0] == "File Path": cur = r[1].split('/')[-1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].isdigit():
        try: samples, inst = int(r[6]), int(r[7])
        except ValueError: continue
        g = lambda n: r[hdr.index(n)]
        agg.append((samples, inst, cur, r[0], r[1].strip()[:80], g("stall_no_inst"), g("stall_wait"), g("stall_long_sb"), g("stall_short_sb"), g("stall_math"), g("Avg. Threads Executed")))
tot = sum(a[0] for a in agg); toti = sum(a[1] for a in agg)
print("total samples", tot, "warp-inst", toti)
for a in sorted(agg, reverse=True)[:top]:
    print(f"{100*a[0]/tot:5.1f}%s {100*a[1]/toti:5.1f}%i noinst={a[5]:>6} wait={a[6]:>6} lsb={a[7]:>6} ssb={a[8]:>6} math={a[9]:>5} thr={a[10]:>3} {a[2]}:{a[3]} {a[4]}")
