"""Golden vectors of Solver::solve_shortchar_order_0 (reference src/solver/solver.tpp:150-235, 1564-1757;
Model::compute_radiation_field_shortchar_order_0, model.cpp:313-349) from the UNMODIFIED compiled reference, for the golden models
(case_<name>.npz holds the inputs): radiation.J, radiation.u (small cases), and the local Lambda operator (one element per
(point, line), the sum of the two half-ray terms of every ray pair and quadrature point).

Note on Lambda: for a half ray without any segment (an origin on the boundary looking outward) the reference evaluates
expm1f(-dtau)/dtau on an UNINITIALISED dtau (solver.tpp:1580,1621); `interior` marks the points whose Lambda values are defined, and
the parity tests compare values there only (counts, J and u everywhere). For three cases also the populations and Jeff after ONE iteration of compute_level_populations_shortchar. Also checks the oracle's restatement against the reference."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import util  # noqa: E402
from magritte_b200 import modelbuild as mb  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from oracle import refcore  # noqa: E402


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = dict(util.small_specs())
    specs["wind3d_adaptive"] = mb.with_adaptive_directions(specs["wind3d_small"])
    for name, spec in specs.items():
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", "sc_" + name), "sc_" + name, spec)
            mb.prepare(m)
            if name == "wind3d_spiral_2spec":
                m.parameters.sum_opacity_emissivity_over_all_lines = False
            m.compute_radiation_field_shortchar_order_0()
        flat, _ = util.load_golden(name)
        N, L = int(flat["npoints"]), int(flat["nlines"])
        out = {"J": np.array(m.radiation.J)}
        u = np.array(m.radiation.u)
        if u.size <= 100000:
            out["u"] = u
        val, cnt = np.zeros((N, L)), np.zeros((N, L), dtype=np.uint32)
        for (p, l), lst in util.lambda_dense(m, L).items():
            cnt[p, l] = len(lst)
            if lst:
                assert lst[0][0] == p and len(lst) == 1
                val[p, l] = lst[0][1]
        out["lambda_val"], out["lambda_cnt"] = val, cnt
        out["interior"] = np.asarray(flat["point2boundary"]) == N
        if name in ("c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"):
            # ONE iteration of Model::compute_level_populations_shortchar (model.cpp:859-953) from the LTE start: the statistical
            # equilibrium of a point reads that point's Lambda element only, so interior points are defined by the reference
            with util.quiet():
                m1 = mb.build(core, core.IoPython("mem", "sc1_" + name), "sc1_" + name, spec)
                mb.prepare(m1)
                if name == "wind3d_spiral_2spec":
                    m1.parameters.sum_opacity_emissivity_over_all_lines = False
                m1.parameters.use_adaptive_Ng_acceleration = False
                assert m1.compute_level_populations_shortchar(False, 1) == 1
            for l, ls in enumerate(m1.lines.lineProducingSpecies):
                out[f"iter1_population_{l}"] = np.array(ls.population, dtype=np.float64)
                out[f"iter1_Jeff_{l}"] = np.array(ls.Jeff)
        np.savez_compressed(os.path.join(HERE, f"shortchar_{name}.npz"), **out)
        o = orc.Oracle(flat, store_u=True)
        Jo = o.shortchar()
        it = out["interior"]
        print(name, "J eq", np.array_equal(Jo, out["J"]), "relerr", util.relerr(Jo, out["J"]), "cnt eq", np.array_equal(o.L_cnt, cnt),
              "lam interior eq", np.array_equal(o.L_val[it, :, 0], val[it]), util.relerr(o.L_val[it, :, 0], val[it]),
              "u eq", np.array_equal(o.u, u))  # u of every case is checked here; only the small ones are stored


if __name__ == "__main__":
    main()
