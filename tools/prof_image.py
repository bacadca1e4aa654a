#!/usr/bin/env python
"""Imaging workload for timing / ncu: configs[4]-shaped image (nx x ny pixels x nfreqs bins around CO J=2-1) of the configs[2] wind model.

python tools/prof_image.py [--npoints 100000] [--nx 512] [--ny 512] [--nfreqs 100] [--reps 2]"""
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
    ap.add_argument("--nx", type=int, default=512)
    ap.add_argument("--ny", type=int, default=512)
    ap.add_argument("--nfreqs", type=int, default=100)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--lattice", type=int, default=0, help="configs[4]: jittered-lattice sphere with about this many points (10000000)")
    ap.add_argument("--check", action="store_true", help="32x32 image of the same model against the oracle (full-size parity)")
    a = ap.parse_args()
    from magritte_b200 import cabi, modelbuild as mb, synth
    import time
    t0 = time.time()
    if a.lattice:
        flat = synth.flat_from_spec(mb.spec_lattice_sphere(npoints=a.lattice))
        nu0 = float(flat["line"][0])
    else:
        flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=a.npoints, nside_rays=1, nquads=3, seed=1234))
        nu0 = float(flat["line"][1])
    print(json.dumps({"model_points": int(flat["npoints"]), "nboundary": int(flat["nboundary"]), "build_s": time.time() - t0}))
    dv = 20.0e3 / 2.99792458e8
    nu = np.linspace(nu0 * (1 - dv), nu0 * (1 + dv), a.nfreqs)
    ctx = cabi.Context(0)
    ctx.set_model(flat, frequencies=False)
    for d in [(1.0, 0.0, 0.0), (0.3, 0.5, 0.81)][:a.reps]:
        X, Y, I, st = ctx.image(2, 3, d, a.nx, a.ny, nu)
        r = st.as_dict()
        r["pixel_freqs_per_s"] = r["n_solves"] / (r["ms_total"] * 1e-3)
        r["I_sum"] = float(I.sum())
        print(json.dumps(r))
    if a.check:
        from oracle import oracle as orc
        f2 = dict(flat)
        f2["nfreqs"] = a.nfreqs
        f2["nu"] = np.ascontiguousarray(np.broadcast_to(nu, (flat["npoints"], a.nfreqs)))
        f2["corresponding_line"] = np.zeros(a.nfreqs, dtype=np.uint32)
        d = (0.3, 0.5, 0.81)
        t0 = time.time()
        want = orc.Oracle(f2, store_u=False).image(d, 32, 32, 3)
        t1 = time.time()
        X, Y, I, st = ctx.image(2, 3, d, 32, 32, nu)
        err = float(np.max(np.abs(I - want["I"]) / np.maximum(np.abs(want["I"]), 1e-300)))
        print(json.dumps({"check": "32x32 image vs oracle", "bit_identical": bool(np.array_equal(I, want["I"])), "max_rel_err": err,
                          "oracle_s": t1 - t0, "gpu_ms": st.ms_total, "ImX_equal": bool(np.array_equal(X, want["ImX"]))}))
        assert err <= 1e-10
    ctx.close()


if __name__ == "__main__":
    main()
