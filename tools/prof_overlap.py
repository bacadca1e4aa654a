#!/usr/bin/env python
"""Does overlapping the pipeline's memory-bound and compute-bound kernels pay? Two contexts (two streams), each on half of the origins,
from two host threads, against one context on all origins. python tools/prof_overlap.py [--npoints 40000] [--nside 2]"""
import argparse, json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser(); ap.add_argument("--npoints", type=int, default=40000); ap.add_argument("--nside", type=int, default=2)
a = ap.parse_args()
from magritte_b200 import cabi, modelbuild as mb, synth
from magritte_b200.flatten import approx_of
flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=a.npoints, nside_rays=a.nside, nquads=11, seed=1234))
approx = approx_of(flat["params"]); N = int(flat["npoints"])
one = cabi.Context(0); one.set_model(flat)
for _ in range(2):
    t0 = time.perf_counter(); one.run(approx); t1 = time.perf_counter()
print(json.dumps({"one_context_s": t1 - t0}))
one.close()
halves = [(0, N // 2), (N // 2, N)]
ctxs = [cabi.Context(0) for _ in halves]
for c in ctxs: c.set_model(flat)
def work(c, rng): c.run(approx, rng)
for rep in range(2):
    th = [threading.Thread(target=work, args=(c, r)) for c, r in zip(ctxs, halves)]
    t0 = time.perf_counter(); [t.start() for t in th]; [t.join() for t in th]; t1 = time.perf_counter()
print(json.dumps({"two_contexts_s": t1 - t0}))
for c, r in zip(ctxs, halves):
    t0 = time.perf_counter(); c.run(approx, r); print(json.dumps({"half_alone_s": time.perf_counter() - t0}))
