"""Golden vectors of the new imager (Solver::image_feautrier_order_2_new_imager) from the UNMODIFIED compiled reference
(oracle/_ref). Run in the build container; the .npz files are committed.

Per case: the model is the one of case_<name>.npz (same flat inputs, asserted here), re-discretised with
Model::compute_spectral_discretisation(nu_min, nu_max, n) around one line, then imaged with
Model::compute_image_new(rx, ry, rz, nx, ny) in several directions. Stored: the image frequencies, and per direction
Image.ImX, Image.ImY and Image.I.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import util  # noqa: E402
from magritte_b200 import flatten, modelbuild as mb  # noqa: E402
from oracle import refcore  # noqa: E402

CASES = {  # name: (dimension, line index, half width [m/s], nfreqs, directions, (nx, ny))
    "wind3d_small": (3, 1, 20e3, 31, [(1, 0, 0), (0.3, 0.5, 0.81), (0, 0, 1), (-0.6, 0.1, -0.4)], (24, 20)),
    "wind3d_spiral_2spec": (3, 2, 70e3, 24, [(0.3, -0.5, 0.2), (0, 1, 0)], (16, 16)),
    "c2_vanzadelhoff": (1, 0, 3e3, 33, [(0, 1, 0)], (24, 24)),
    "velgrad_1d": (1, 0, 40e3, 40, [(0, 1, 0)], (12, 12)),
    "c1_all_constant": (1, 0, 1e3, 7, [(0, 1, 0)], (4, 4)),
}


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = util.small_specs()
    for name, (dim, li, dv, nf, dirs, (nx, ny)) in CASES.items():
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", name), name, specs[name])
            mb.prepare(m)
            if name == "wind3d_spiral_2spec":
                m.parameters.sum_opacity_emissivity_over_all_lines = False
        nu0 = float(np.array(m.lines.line)[li])
        with util.quiet():
            m.compute_spectral_discretisation(nu0 * (1 - dv / 2.99792458e8), nu0 * (1 + dv / 2.99792458e8), nf)
        flat = flatten.flatten_model(m)
        case, _ = util.load_golden(name)
        for k in ("position", "velocity", "neighbors", "emissivity", "opacity", "inverse_width", "population_tot", "point2boundary"):
            assert np.array_equal(flat[k], case[k]), k
        nu = flat["nu"]
        assert (nu == nu[0]).all()
        out = {"nu": np.ascontiguousarray(nu[0]), "dimension": np.array(dim), "nx": np.array(nx), "ny": np.array(ny),
               "directions": np.array(dirs, dtype=np.float64)}
        for i, d in enumerate(dirs):
            with util.quiet():
                m.compute_image_new(float(d[0]), float(d[1]), float(d[2]), nx, ny)
            im = m.images[-1]
            out[f"ImX_{i}"], out[f"ImY_{i}"], out[f"I_{i}"] = np.array(im.ImX), np.array(im.ImY), np.array(im.I)
        path = os.path.join(HERE, f"image_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "dirs", len(dirs), "I", out["I_0"].shape, "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
