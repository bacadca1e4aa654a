"""Parity at the sizes BASELINE.json names, inside the driver-run GPU suite (round 1 had these only in bench.py and in builder logs):

  * configs[2] in full (112 288 points, 96 ray pairs, 440 bins, one radiation-field call): >= 512 oracle rows (reference elimination
    order) against the same rows of the CUDA result, <= 1e-10; the work counters against the model's own counts;
  * EVERY ray path of the five golden models -- all (origin, ray pair, line) -- bit for bit against the oracle's tracer, which
    tests/test_oracle_golden.py pins to the compiled reference's scratch arrays (the golden files hold 40 sampled paths per case);
  * configs[3]'s shape (spiral wind, CO + SiO-like, 880 bins): K memory-sparse level-population iterations on the device against a loop
    driven by the ORACLE's radiation field and the host twin's long-double statistical equilibrium.
MGB_FULLSIZE=0 skips the two model-building tests (about 2 minutes together)."""
import os

import numpy as np
import pytest

import util
from test_host_model import assert_populations_close

pytestmark = pytest.mark.gpu
FULL = os.environ.get("MGB_FULLSIZE", "1") != "0"


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_every_ray_path_bit_exact(name):
    from magritte_b200 import cabi
    from oracle import oracle as orc
    flat, ref = util.load_golden(name)
    o = orc.Oracle(flat, store_u=False)
    cap = max(8, o.max_length())
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        n = 0
        for org in range(int(flat["npoints"])):
            for rr in range(int(flat["hnrays"])):
                for l in range(int(flat["nlines"])):
                    util.assert_trace_equal(ctx.trace(org, rr, l, cap=2 * cap + 1), o.trace(org, rr, l, cap=cap), f"{name} ({org},{rr},{l})")
                    n += 1
    finally:
        ctx.close()
    assert n == int(flat["npoints"]) * int(flat["hnrays"]) * int(flat["nlines"])


@pytest.mark.skipif(not FULL, reason="MGB_FULLSIZE=0")
def test_config2_full_size_rows_against_oracle():
    from magritte_b200 import cabi, modelbuild as mb, synth
    from magritte_b200.flatten import approx_of
    from oracle import oracle as orc
    flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=100000, nside_rays=4, nquads=11, seed=1234))
    N, H, F = int(flat["npoints"]), int(flat["hnrays"]), int(flat["nfreqs"])
    assert (N, H, F) == (112288, 96, 440)
    approx = approx_of(flat["params"])
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        st = ctx.run(approx)
        J = ctx.get_J()
        lam, cnt = ctx.get_lambda()
    finally:
        ctx.close()
    assert st.n_solves == N * H * F
    o = orc.Oracle(flat, approx=approx, store_u=False, threads=os.cpu_count() or 1)
    worst = worst_l = 0.0
    rows = 0
    for start, count in ((1000, 176), (N // 2, 176), (N - 12288 - 176, 176), (N - 64, 64)):  # inner wind, middle, outermost interior, boundary shell
        o.solve(o_range=(start, start + count), rr_range=(0, H), half_length=4096)
        worst = max(worst, util.relerr(J[start:start + count], o.J[start:start + count]))
        worst_l = max(worst_l, util.relerr(lam[start:start + count], o.L_val[start:start + count, :, 0]))
        assert np.array_equal(cnt[start:start + count], o.L_cnt[start:start + count])
        rows += count
    print(f"configs[2] full size: {rows} oracle rows, max rel err J {worst:.2e}, Lambda {worst_l:.2e}")
    assert rows >= 512 and worst <= 1e-10 and worst_l <= 1e-10


@pytest.mark.skipif(not FULL, reason="MGB_FULLSIZE=0")
def test_config3_shape_sparse_levelpop_against_oracle_driven_loop():
    from magritte_b200 import model as hostmodel, modelbuild as mb, synth
    from oracle import oracle as orc
    K, pairs = 2, 2
    spec = mb.spec_wind_3d(npoints=int(os.environ.get("MGB_C3_LP_POINTS", "12000")), nside_rays=8, nquads=11, spiral=True, species=("co", "sio"), seed=1234)
    h = len(spec["direction"]) // 2
    idx = [0, h // 2, h, h + h // 2][:2 * pairs] if pairs == 2 else None
    spec["direction"], spec["weight"] = np.ascontiguousarray(spec["direction"][idx]), np.ascontiguousarray(spec["weight"][idx])
    geom = synth.flat_from_spec(spec, with_state=False)
    N, L, nq = int(geom["npoints"]), int(geom["nlines"]), int(geom["nquads"])
    models = []
    for dev in (True, False):
        m = hostmodel.from_spec(spec, geom, 0, use_adaptive_Ng_acceleration=False)
        m.device_level_populations = dev
        m.compute_spectral_discretisation(); m.compute_inverse_line_widths(); m.compute_LTE_level_populations()
        models.append(m)
    md, mh = models
    assert md.compute_level_populations_sparse(False, K) == K
    # the oracle-driven loop: radiation field on the CPU in the reference's elimination order, Jeff and statistical equilibrium by the host twin
    w = np.repeat(np.asarray(geom["quad_weights"]), np.asarray(geom["nrad"]), axis=0)
    f2 = dict(geom)
    f2["nfreqs"] = L * nq
    f2["nu"] = mh.get_nu()
    f2["corresponding_line"] = (np.arange(L * nq, dtype=np.uint32) // nq).astype(np.uint32)
    f2["inverse_width"] = mh.get_inverse_width()
    f2["population_tot"] = np.ascontiguousarray(np.stack([mh.get_population_tot(l) for l in range(int(geom["nlspecs"]))]))
    for _ in range(K):
        f2["emissivity"], f2["opacity"] = mh.get_emissivity(), mh.get_opacity()
        o = orc.Oracle(f2, store_u=False, threads=os.cpu_count() or 1)
        o.solve(half_length=8192)
        Jl = (o.J.reshape(N, L, nq) * w[None]).sum(axis=2)
        mh.set_radiation_field(o.J, np.ascontiguousarray(o.L_val[:, :, 0]), o.L_cnt, Jl)
        mh.compute_Jeff_sparse()
        mh.compute_level_populations_from_stateq_sparse()
    for l in range(int(geom["nlspecs"])):
        assert_populations_close(md.get_population(l), mh.get_population(l), md.get_population_tot(l), f"species {l}", abs_tot=1e-13)
        # radiation quantities: only where both levels of the line hold a share of the molecules that the statistical-equilibrium solve
        # determines well (> 1e-3 of the total: the populations' 1e-13-of-the-total floor is then 1e-10 of the level). Lines between emptier
        # levels (J > 15 in the cold outer wind) have emissivities made of solver noise on BOTH sides and a J of 1e-29 that nothing reads.
        ld = spec["linedata"][l]
        pop, tot = mh.get_population(l), mh.get_population_tot(l)[:, None]
        solid = (pop[:, np.asarray(ld["irad"])] > 1e-3 * tot) & (pop[:, np.asarray(ld["jrad"])] > 1e-3 * tot)
        assert solid.sum() > 3 * N
        # ... and even there J integrates emissivities of OTHER points along the rays, where the same level may hold 1e-6 of the molecules
        # and carry 1e-13 / 1e-6 = 1e-7 of relative solver noise: the bound is 1e-6, the worst case is printed
        for what, a_, b_ in (("Jlin", md.get_Jlin(l), mh.get_Jlin(l)), ("Jeff", md.get_Jeff(l), mh.get_Jeff(l))):
            worst = float(np.max(np.abs(a_ - b_)[solid] / np.abs(b_)[solid]))
            print(f"species {l} {what}: worst relative difference on {int(solid.sum())} well-populated (point, line) pairs {worst:.2e}")
            assert worst <= 1e-6
