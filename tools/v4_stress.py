#!/usr/bin/env python
"""Repeat the streaming pipeline (MGB_SOLVE_KERNEL=4) against k_solve on the golden models and report any bit that differs."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util
from magritte_b200 import cabi
from magritte_b200.flatten import approx_of

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for name in util.GOLDEN_CASES + ["wind3d_adaptive"]:
    flat, _ = util.load_golden(name)
    ctx = cabi.Context(0)
    ctx.set_model(flat)
    ctx.set_params(store_intensities=1)
    for approx in (2, 1, 0):
        os.environ["MGB_SOLVE_KERNEL"] = "1"
        ctx.run(approx); J1 = ctx.get_J().copy(); u1 = ctx.get_u().copy(); l1, c1 = ctx.get_lambda()
        os.environ["MGB_SOLVE_KERNEL"] = "4"
        bad = 0
        for i in range(reps):
            ctx.run(approx); J = ctx.get_J(); u = ctx.get_u(); l4, c4 = ctx.get_lambda()
            if not (np.array_equal(u, u1) and np.array_equal(l4, l1)):
                bad += 1
                d = np.argwhere(u != u1)
                print(name, approx, "rep", i, "u differs at", len(d), "places; first", d[:5].tolist(), "rr,o,f; values", [(u[tuple(x)], u1[tuple(x)]) for x in d[:3]], flush=True)
                dl = np.argwhere(l4 != l1)
                print("   lambda differs at", len(dl), dl[:5].tolist())
        print(name, "approx", approx, "bad", bad, "of", reps, flush=True)
    ctx.close()
