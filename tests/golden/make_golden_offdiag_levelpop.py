"""Golden level populations with a BANDED approximate Lambda operator (parameters.n_off_diag = K > 0: the rate equations of
neighbouring points couple, LineProducingSpecies::update_using_statistical_equilibrium lineProducingSpecies.tpp:377-581 solves one
global sparse system) from the UNMODIFIED compiled reference (oracle/_ref). Per case, with the model and K of offdiag_<name>.npz:
  * one iteration of Model::compute_level_populations(False, 1): Jeff, populations, error_max / error_mean -- the statistical-
    equilibrium step on the J and the Lambda lists that offdiag_<name>.npz already holds;
  * K2 iterations without Ng acceleration: populations (the GPU flow's target)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import util  # noqa: E402
from magritte_b200 import modelbuild as mb  # noqa: E402
from oracle import refcore  # noqa: E402

CASES = {"c2_vanzadelhoff": (2, 3), "wind3d_small": (2, 3), "c1_all_constant": (1, 2), "wind3d_spiral_2spec": (1, 2)}


def fresh(core, name, spec, K):
    with util.quiet():
        m = mb.build(core, core.IoPython("mem", name), name, spec)
        mb.prepare(m)
        if name == "wind3d_spiral_2spec":
            m.parameters.sum_opacity_emissivity_over_all_lines = False
        m.parameters.n_off_diag = K
        m.parameters.use_adaptive_Ng_acceleration = False
    return m


def pops(m):
    return [np.array(l.population, dtype=np.float64) for l in m.lines.lineProducingSpecies]


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = util.small_specs()
    for name, (K, K2) in CASES.items():
        m1 = fresh(core, name, specs[name], K)
        with util.quiet():
            m1.compute_level_populations(False, 1)
        out = {"n_off_diag": np.array(K), "iterations": np.array(K2), "nspecies": np.array(len(m1.lines.lineProducingSpecies)),
               "iter1_error_max": np.array(m1.error_max), "iter1_error_mean": np.array(m1.error_mean)}
        for l, (P, ls) in enumerate(zip(pops(m1), m1.lines.lineProducingSpecies)):
            out[f"iter1_population_{l}"] = P
            out[f"iter1_Jeff_{l}"] = np.array(ls.Jeff)
        m2 = fresh(core, name, specs[name], K)
        with util.quiet():
            it = m2.compute_level_populations(False, K2)
        assert it == K2
        for l, P in enumerate(pops(m2)):
            out[f"final_population_{l}"] = P
        out["final_error_max"] = np.array(m2.error_max)
        path = os.path.join(HERE, f"offdiag_levelpop_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "K", K, "iterations", K2, "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
