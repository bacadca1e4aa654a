"""Golden vectors of Solver::solve_feautrier_order_2_anis (reference src/solver/solver.tpp:365-456;
Model::compute_radiation_field_feautrier_order_2_anis, model.cpp:435-470) from the UNMODIFIED compiled reference: lspec.J and the
anisotropic tensor elements J2_0, J2_1_Re, J2_1_Im, J2_2_Re, J2_2_Im per species for the golden models (case_<name>.npz holds the inputs)."""
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


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = dict(util.small_specs())
    specs["wind3d_adaptive"] = mb.with_adaptive_directions(specs["wind3d_small"])
    for name, spec in specs.items():
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", "anis_" + name), "anis_" + name, spec)
            mb.prepare(m)
            if name == "wind3d_spiral_2spec":
                m.parameters.sum_opacity_emissivity_over_all_lines = False
            m.compute_radiation_field_feautrier_order_2_anis()
        out = {}
        for q in ("J", "J2_0", "J2_1_Re", "J2_1_Im", "J2_2_Re", "J2_2_Im"):
            out[q] = np.concatenate([np.array(getattr(ls, q)) for ls in m.lines.lineProducingSpecies], axis=1)
        np.savez_compressed(os.path.join(HERE, f"anis_{name}.npz"), **out)
        print(name, out["J"].shape, float(np.abs(out["J2_0"]).max()))


if __name__ == "__main__":
    main()
