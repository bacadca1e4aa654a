"""Golden vectors of the memory-sparse solver family from the UNMODIFIED compiled reference (oracle/_ref):
Model::compute_radiation_field_feautrier_order_2_sparse (model.cpp:474-510, Solver::solve_feautrier_order_2_sparse solver.tpp:297-363),
compute_Jeff_sparse (:549-570) and compute_level_populations_sparse (:756-850). Run in the build container; the .pkl files are committed.

Per case, from the LTE start of levelpop_<name>.pkl (same model, same parameters):
  * lspec.J [npoints][nrad] of every species after one sparse radiation-field call (compute_Jeff_sparse is not bound by the
    reference's pybind module; it is covered through the loop below);
  * populations and Jeff after K iterations of compute_level_populations_sparse WITHOUT Ng acceleration;
  * iteration count and final populations of the default (adaptive Ng) run.
"""
import os
import pickle
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)

import util  # noqa: E402
from make_golden_levelpop import CASES, fresh, pops  # noqa: E402
from oracle import refcore  # noqa: E402


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = util.small_specs()
    for name, K in CASES.items():
        m = fresh(core, name, specs[name])
        with util.quiet():
            m.compute_radiation_field_feautrier_order_2_sparse()
        out = {"K": K, "J_species": [np.array(l.J) for l in m.lines.lineProducingSpecies]}
        m1 = fresh(core, name, specs[name])
        with util.quiet():
            m1.parameters.use_adaptive_Ng_acceleration = False
            it = m1.compute_level_populations_sparse(False, K)
        assert it == K
        out["noNg_population"] = pops(m1)
        out["noNg_Jeff"] = [np.array(l.Jeff) for l in m1.lines.lineProducingSpecies]
        out["noNg_error_max"] = list(m1.error_max)
        m2 = fresh(core, name, specs[name])
        with util.quiet():
            it2 = m2.compute_level_populations_sparse(True, 60)
        out["Ng_iterations"], out["Ng_population"], out["Ng_error_max"] = it2, pops(m2), list(m2.error_max)
        path = os.path.join(HERE, f"sparse_{name}.pkl")
        with open(path, "wb") as f:
            pickle.dump(out, f, protocol=4)
        print(name, "K", K, "Ng iterations", it2, "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
