"""configs[3]-shaped workload: spiral-perturbed 3D wind, two line-producing species (CO and the SiO-like species
derived from it: 80 lines x 11 quadrature points = 880 frequency bins), HEALPix nside 8 ray set restricted to a ray-pair subset
(the work of one (o, rr, f) solve does not depend on how many pairs are run -- the same device the reference's CPU baseline uses).

Default size keeps the GPU suite short; MGB_C3_POINTS=1000000 runs the point count BASELINE.json names (about 3 min of host-side
Delaunay + flattening, 10 GB of host arrays). The checks are the ones that do not depend on the size:
  * oracle rows (reference elimination order) of a block of origins against the same rows of the CUDA result, <= 1e-10;
  * the number of solves and of ray entries the library reports against the model's own counts;
  * a second pass over cached ray paths reproduces J bit for bit.
"""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-10
NPOINTS = int(os.environ.get("MGB_C3_POINTS", "40000"))
PAIRS = int(os.environ.get("MGB_C3_PAIRS", "2"))
SAMPLE = int(os.environ.get("MGB_C3_SAMPLE", "1024"))


def test_config3_sample_against_oracle():
    from magritte_b200 import cabi, modelbuild as mb, synth
    from magritte_b200.flatten import approx_of
    from oracle import oracle as orc
    spec = mb.spec_wind_3d(npoints=NPOINTS, nside_rays=8, nquads=11, spiral=True, species=("co", "sio"), seed=1234)
    h = len(spec["direction"]) // 2
    d, w = spec["direction"], spec["weight"]
    spec["direction"] = np.concatenate([d[:PAIRS], d[h:h + PAIRS]], axis=0)
    spec["weight"] = np.concatenate([w[:PAIRS], w[h:h + PAIRS]])
    flat = synth.flat_from_spec(spec)
    N, H, F = int(flat["npoints"]), int(flat["hnrays"]), int(flat["nfreqs"])
    assert (H, F) == (PAIRS, 880) and int(flat["nlspecs"]) == 2
    approx = approx_of(flat["params"])
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        st = ctx.run(approx)
        assert st.n_solves == N * H * F
        J = ctx.get_J()
        st2 = ctx.run(approx)
        assert st2.paths_cached == 1
        assert np.array_equal(ctx.get_J(), J)
    finally:
        ctx.close()
    assert np.all(np.isfinite(J)) and np.all(J > 0.0)
    # a block of origins in the dense inner wind and one in the outer shell region
    cores = os.cpu_count() or 1
    o = orc.Oracle(flat, approx=approx, store_u=False, threads=cores)
    worst = 0.0
    for start in (N // 7, max(0, NPOINTS - SAMPLE)):
        stop = min(N, start + SAMPLE)
        o.solve(o_range=(start, stop), rr_range=(0, H), half_length=8192)
        a, b = J[start:stop], o.J[start:stop]
        worst = max(worst, float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))))
    print(json.dumps({"config3": {"npoints": N, "hnrays": H, "nfreqs": F, "n_solves": int(st.n_solves), "n_entries": int(st.n_entries),
                                  "ms_total": st.ms_total, "ms_total_cached_paths": st2.ms_total,
                                  "solves_per_s": st.n_solves / (st.ms_total * 1e-3), "oracle_origins": 2 * SAMPLE,
                                  "max_rel_err_J": worst}}))
    assert worst <= TOL
