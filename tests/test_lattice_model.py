"""The configs[4] point-cloud generator (magritte_b200.modelbuild.spec_lattice_sphere): structural invariants the tracer relies on,
and a small image of it through the oracle (host only)."""
import numpy as np

from magritte_b200 import modelbuild as mb, synth


def test_lattice_sphere_invariants():
    spec = mb.spec_lattice_sphere(npoints=4000)
    N = len(spec["position"])
    n_nbs, nbs = spec["n_nbs"], spec["nbs"]
    assert 0.8 * 4000 < N < 1.2 * 4000 and n_nbs.sum() == len(nbs) and n_nbs.max() == 26
    cum = np.concatenate([[0], np.cumsum(n_nbs, dtype=np.int64)]).astype(np.int64)
    pairs = set()
    for p in range(N):
        for q in nbs[cum[p]:cum[p + 1]]:
            pairs.add((p, int(q)))
    assert all((q, p) in pairs for (p, q) in pairs)  # symmetric neighbour lists
    b = set(spec["boundary2point"].tolist())
    assert b == set(np.nonzero(n_nbs < 26)[0].tolist()) and len(b) > 0
    r = np.linalg.norm(spec["position"], axis=1)
    assert r.max() < 1.05e14
    # every interior point is fully surrounded: a ray cannot leave the cloud without passing a boundary point's cell
    assert (n_nbs[np.setdiff1d(np.arange(N), spec["boundary2point"])] == 26).all()


def test_lattice_sphere_image_through_oracle():
    from oracle import oracle as orc
    flat = synth.flat_from_spec(mb.spec_lattice_sphere(npoints=3000))
    nu0, dv = float(flat["line"][0]), 20e3 / 2.99792458e8
    nu = np.linspace(nu0 * (1 - dv), nu0 * (1 + dv), 7)
    flat["nfreqs"] = 7
    flat["nu"] = np.ascontiguousarray(np.broadcast_to(nu, (flat["npoints"], 7)))
    flat["corresponding_line"] = np.zeros(7, dtype=np.uint32)
    r = orc.Oracle(flat, store_u=False).image((0.3, 0.5, 0.81), 12, 12, 3)
    assert np.isfinite(r["I"]).all() and (r["I"] > 0).all() and r["n_tot"].max() > 5
