#!/usr/bin/env python
"""Small fixed workload for ncu captures of the solve kernel (profiling aid; not a bench).

python tools/prof_solve.py [--npoints 8000] [--nside 2] [--reps 2]
Builds the configs[2]-shaped wind model at a reduced size, runs the Feautrier order-2 pass `reps` times through the
C ABI and prints the library's own event timings.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--npoints", type=int, default=8000)
    ap.add_argument("--nside", type=int, default=2)
    ap.add_argument("--nquads", type=int, default=11)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--approx", type=int, default=-1)
    a = ap.parse_args()
    from magritte_b200 import cabi, modelbuild as mb, synth
    from magritte_b200.flatten import approx_of
    flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=a.npoints, nside_rays=a.nside, nquads=a.nquads, seed=1234))
    approx = approx_of(flat["params"]) if a.approx < 0 else a.approx
    ctx = cabi.Context(0)
    ctx.set_model(flat)
    for _ in range(a.reps):
        st = ctx.run(approx)
        d = st.as_dict()
        d["solves_per_s"] = d["n_solves"] / (d["ms_total"] * 1e-3)
        print(json.dumps(d))
    ctx.close()


if __name__ == "__main__":
    main()
