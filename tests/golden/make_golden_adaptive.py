"""Golden vectors for Rays::use_adaptive_directions (a ray set per point; reference src/model/geometry/rays/rays.hpp:23-31,
rays.cpp:37-100, selected at src/model/model.cpp:357-389) from the UNMODIFIED compiled reference (oracle/_ref): the small 3D wind
of the other goldens with every point's ray set rotated and re-weighted (modelbuild.with_adaptive_directions). The file holds the
flat inputs and the reference's J, Lambda lists, Model.column, and K level-population iterations without Ng acceleration."""
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

K = 3


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    spec = mb.with_adaptive_directions(util.small_specs()["wind3d_small"])
    name = "wind3d_adaptive"
    with util.quiet():
        m = mb.build(core, core.IoPython("mem", name), name, spec)
        mb.prepare(m)
        assert m.geometry.rays.use_adaptive_directions
        m.compute_radiation_field_feautrier_order_2()
    flat = flatten.flatten_model(m)
    out = {}
    for k, v in flat.items():
        if k == "params":
            for pk, pv in v.items():
                out["param_" + pk] = np.array(pv)
        else:
            out[k] = np.asarray(v)
    out["ref_J"] = np.array(m.radiation.J)
    lam = util.lambda_dense(m, flat["nlines"])
    val = np.zeros((flat["npoints"], flat["nlines"]))
    cnt = np.zeros((flat["npoints"], flat["nlines"]), dtype=np.uint32)
    for (p, l), lst in lam.items():
        cnt[p, l] = len(lst)
        if lst:
            assert lst[0][0] == p and len(lst) == 1
            val[p, l] = lst[0][1]
    out["ref_lambda_val"], out["ref_lambda_cnt"] = val, cnt
    rho = np.asarray(spec["abundance"])[:, 1].copy()
    with util.quiet():
        m.density.set(rho)
        m.set_column()
    out["ref_density"], out["ref_column"] = rho, np.array(m.column)
    np.savez_compressed(os.path.join(HERE, f"case_{name}.npz"), **out)
    # level populations: K iterations, no Ng
    lp = dict(species=hostmodel.linedata_of(m), parameters=hostmodel.parameters_of(m), abundance=np.array(m.chemistry.species.abundance), K=K)
    m.parameters.use_adaptive_Ng_acceleration = False
    with util.quiet():
        assert m.compute_level_populations(False, K) == K
    lp["noNg_population"] = [np.array(ls.population, dtype=np.float64) for ls in m.lines.lineProducingSpecies]
    lp["noNg_Jeff"] = [np.array(ls.Jeff) for ls in m.lines.lineProducingSpecies]
    lp["noNg_J"] = np.array(m.radiation.J)[::7]
    lp["noNg_error_max"] = list(m.error_max)
    with open(os.path.join(HERE, f"levelpop_{name}.pkl"), "wb") as f:
        pickle.dump(lp, f)
    print(name, out["ref_J"].shape, "error_max", lp["noNg_error_max"])


if __name__ == "__main__":
    main()
