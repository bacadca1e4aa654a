"""Golden vectors of the level-population loop (Model::compute_level_populations, reference model.cpp:655-748) from the
UNMODIFIED compiled reference (oracle/_ref). Run in the build container; the .npz files are committed.

Per case: the line data / abundances / parameters the loop reads (magritte_b200.model.linedata_of), the LTE start, and
  * after 1 and after K iterations WITHOUT Ng acceleration: populations, Jeff, a sample of J (the 1e-10 parity input of SURVEY.md 8c);
  * the full run WITH the default (adaptive) Ng acceleration: iteration count and final populations.
"""
import os
import pickle
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import util  # noqa: E402
from magritte_b200 import flatten, model as hostmodel, modelbuild as mb  # noqa: E402
from oracle import refcore  # noqa: E402

CASES = {"c2_vanzadelhoff": 4, "wind3d_small": 3, "wind3d_spiral_2spec": 3, "c1_all_constant": 3}


def fresh(core, name, spec):
    with util.quiet():
        m = mb.build(core, core.IoPython("mem", name), name, spec)
        mb.prepare(m)
        if name == "wind3d_spiral_2spec":
            m.parameters.sum_opacity_emissivity_over_all_lines = False
    return m


def pops(m):
    return [np.array(l.population, dtype=np.float64) for l in m.lines.lineProducingSpecies]


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = util.small_specs()
    for name, K in CASES.items():
        m = fresh(core, name, specs[name])
        out = {"species": hostmodel.linedata_of(m), "parameters": hostmodel.parameters_of(m),
               "abundance": np.array(m.chemistry.species.abundance, dtype=np.float64), "K": K, "lte_population": pops(m),
               "lte_emissivity": np.array(m.lines.emissivity), "lte_opacity": np.array(m.lines.opacity)}
        m1 = fresh(core, name, specs[name])  # one iteration: the statistical-equilibrium step on the J / Lambda of case_<name>.npz
        with util.quiet():
            m1.parameters.use_adaptive_Ng_acceleration = False
            m1.compute_level_populations(False, 1)
        out["iter1_population"] = pops(m1)
        out["iter1_Jeff"] = [np.array(l.Jeff) for l in m1.lines.lineProducingSpecies]
        out["iter1_error_max"], out["iter1_error_mean"] = list(m1.error_max), list(m1.error_mean)
        with util.quiet():
            m.parameters.use_adaptive_Ng_acceleration = False
            it = m.compute_level_populations(False, K)
        assert it == K
        out["noNg_population"] = pops(m)
        out["noNg_Jeff"] = [np.array(l.Jeff) for l in m.lines.lineProducingSpecies]
        out["noNg_J"] = np.array(m.radiation.J)[::7]  # every 7th point
        out["noNg_error_max"] = list(m.error_max)
        # default (adaptive Ng) run to convergence on a fresh model
        m2 = fresh(core, name, specs[name])
        with util.quiet():
            it2 = m2.compute_level_populations(True, 60)
        out["Ng_iterations"], out["Ng_population"], out["Ng_error_max"] = it2, pops(m2), list(m2.error_max)
        path = os.path.join(HERE, f"levelpop_{name}.pkl")
        with open(path, "wb") as f:
            pickle.dump(out, f, protocol=4)
        print(name, "K", K, "Ng iterations", it2, "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
