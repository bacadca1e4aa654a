"""Golden vectors of the optical-depth imager (Model::compute_image_optical_depth_new -> Solver::image_optical_depth_new_imager,
solver.tpp:688-733, 2718-2760) from the UNMODIFIED compiled reference (oracle/_ref). Run in the build container; the .npz files are
committed. Same models, frequency bins, directions and pixel counts as image_<name>.npz (make_golden_image.py); stored per
direction: Image.I (the optical depth per pixel and frequency bin).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)

import util  # noqa: E402
from make_golden_image import CASES  # noqa: E402
from magritte_b200 import modelbuild as mb  # noqa: E402
from oracle import refcore  # noqa: E402


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
        out = {}
        for i, d in enumerate(dirs):
            with util.quiet():
                m.compute_image_optical_depth_new(float(d[0]), float(d[1]), float(d[2]), nx, ny)
            im = m.images[-1]
            out[f"ImX_{i}"], out[f"ImY_{i}"], out[f"tau_{i}"] = np.array(im.ImX), np.array(im.ImY), np.array(im.I)
        path = os.path.join(HERE, f"imagetau_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "dirs", len(dirs), "tau", out["tau_0"].shape, "max", float(out["tau_0"].max()), "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
