#!/usr/bin/env python
"""Column-density workload for timing / ncu: Model::set_column on the configs[2]-shaped wind point cloud (HEALPix nside 4).

python tools/prof_column.py [--npoints 100000] [--nside 4] [--reps 3]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--npoints", type=int, default=100000)
    ap.add_argument("--nside", type=int, default=4)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    from magritte_b200 import cabi, modelbuild as mb, synth
    flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=a.npoints, nside_rays=a.nside, nquads=3, nlines_keep=range(2), seed=4321))
    N, R = int(flat["npoints"]), int(flat["nrays"])
    density = 10.0 ** np.random.default_rng(99).uniform(6.0, 14.0, size=N)
    ctx = cabi.Context(0)
    ctx.set_model(flat)
    for _ in range(a.reps):
        col, ms = ctx.set_column(density, with_time=True)
        print(json.dumps({"npoints": N, "nrays": R, "kernel_ms": ms, "half_rays_per_s": N * R / (ms * 1e-3), "sum": float(col.sum())}))
    ctx.close()


if __name__ == "__main__":
    main()
