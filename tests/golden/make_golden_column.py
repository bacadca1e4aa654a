"""Golden vectors of Model::set_column (model.cpp:1143-1154 -> Solver::set_column / get_column, solver.tpp:2953-2991) from the
UNMODIFIED compiled reference (oracle/_ref). Run in the build container; the .npz files are committed.

Per case (the models of case_<name>.npz): a seeded log-uniform `Model::density` [npoints] and the resulting `Model::column`
[nrays][npoints].
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


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = util.small_specs()
    for i, name in enumerate(util.GOLDEN_CASES):
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", name), name, specs[name])
            mb.prepare(m)
        flat = flatten.flatten_model(m)
        case, _ = util.load_golden(name)
        for k in ("position", "neighbors", "point2boundary"):
            assert np.array_equal(flat[k], case[k]), k
        rng = np.random.default_rng(1234 + i)
        density = 10.0 ** rng.uniform(6.0, 14.0, size=flat["npoints"])
        m.density.set(np.ascontiguousarray(density))
        with util.quiet():
            m.set_column()
        column = np.array(m.column)
        assert column.shape == (flat["nrays"], flat["npoints"]), column.shape
        path = os.path.join(HERE, f"column_{name}.npz")
        np.savez_compressed(path, density=density, column=column)
        print(name, column.shape, "max", float(column.max()), "zeros", int((column == 0).sum()), "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
