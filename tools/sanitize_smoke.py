#!/usr/bin/env python
"""Small end-to-end pass for compute-sanitizer (memcheck / racecheck): radiation field (all three ApproximationTypes, banded Lambda),
device level populations, image, column density on the small golden models. Not a test of values (tests/ does that)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from magritte_b200 import cabi  # noqa: E402
from test_host_model import load_levelpop  # noqa: E402

for name in ("c2_vanzadelhoff", "wind3d_spiral_2spec", "wind3d_adaptive"):
    flat, ref = util.load_golden(name)
    ctx = cabi.Context(0)
    ctx.set_model(flat)
    for approx in (2, 1, 0):
        ctx.run(approx)
    J = ctx.get_J()
    if name != "wind3d_adaptive":
        ctx.set_params(n_off_diag=2)
        ctx.run(2)
        ctx.get_lambda_lists()
        ctx.set_params(n_off_diag=0)
        lp = load_levelpop(name)
        ctx.run(2)
        ctx.lp_setup(lp["species"], lp["abundance"], lp["parameters"])
        for l, pop in enumerate(lp["lte_population"]):
            ctx.lp_set_populations(l, np.asarray(pop, dtype=np.float64).reshape(int(flat["npoints"]), -1), flat["population_tot"][l])
        ctx.lp_compute_Jeff()
        ctx.lp_update_statistical_equilibrium()
    if not int(flat["spherical_symmetry"]):
        ctx.set_column(np.ones(int(flat["npoints"])))
        nu = np.linspace(float(flat["line"][0]) * (1 - 7e-5), float(flat["line"][0]) * (1 + 7e-5), 9)
        ctx.image(2, 3, (0.3, 0.5, 0.81), 12, 10, nu)
    ctx.close()
    print(name, "ok", float(J.sum()))
