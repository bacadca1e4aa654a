"""Short characteristics, order 0 (reference Solver::solve_shortchar_order_0, solver.tpp:150-235, 1564-1757): the oracle's restatement
against the compiled reference's golden vectors (tests/golden/shortchar_*.npz, made by make_golden_shortchar.py), and the CUDA path
(mgb_run_shortchar) against the oracle and the goldens.

Bit-exact everywhere except the Lambda VALUES of boundary points: for a half ray without a segment the reference evaluates
expm1f(-dtau)/dtau on an uninitialised dtau (solver.tpp:1580, 1621), so those values are whatever the stack held; the oracle and the
device define that factor as 0 and the comparison covers interior points (counts, J and u are compared everywhere)."""
import os

import numpy as np
import pytest

import util

CASES = util.GOLDEN_CASES + ["wind3d_adaptive"]


def load_shortchar(name):
    return dict(np.load(os.path.join(util.GOLDEN, f"shortchar_{name}.npz")))


@pytest.mark.parametrize("name", CASES)
def test_oracle_shortchar_matches_reference(name):
    from oracle import oracle as orc
    flat, _ = util.load_golden(name)
    g = load_shortchar(name)
    o = orc.Oracle(flat, store_u=True)
    J = o.shortchar()
    assert np.array_equal(J, g["J"])
    assert np.array_equal(o.L_cnt, g["lambda_cnt"])
    it = g["interior"]
    assert it.any()
    assert np.array_equal(o.L_val[it, :, 0], g["lambda_val"][it])
    if "u" in g:
        assert np.array_equal(o.u, g["u"])


def test_oracle_shortchar_restricted_range_is_a_slice():
    """Origins are independent: a restricted (o, rr) range reproduces the rows of the full run's partial sums."""
    from oracle import oracle as orc
    flat, _ = util.load_golden("wind3d_small")
    full = orc.Oracle(flat, store_u=True)
    full.shortchar()
    part = orc.Oracle(flat, store_u=True)
    part.shortchar(o_range=(100, 164))
    assert np.array_equal(part.J[100:164], full.J[100:164])
    assert np.array_equal(part.u[:, 100:164], full.u[:, 100:164])
    assert not part.J[:100].any() and not part.J[164:].any()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_shortchar_against_oracle_and_reference(name):
    from magritte_b200 import cabi
    from magritte_b200.flatten import approx_of
    from oracle import oracle as orc
    flat, _ = util.load_golden(name)
    g = load_shortchar(name)
    o = orc.Oracle(flat, store_u=True)
    o.shortchar()
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        st = ctx.run_shortchar(approx_of(flat["params"]), store_u=True)
        J, u = ctx.get_J(), ctx.get_u()
        val, cnt = ctx.get_lambda()
    finally:
        ctx.close()
    assert st.n_launches > 0
    assert np.array_equal(J, g["J"]) and np.array_equal(J, o.J)
    assert np.array_equal(u, o.u)
    assert np.array_equal(cnt, g["lambda_cnt"])
    assert np.array_equal(val, o.L_val[:, :, 0])  # the oracle's definition on the boundary too
    it = g["interior"]
    assert np.array_equal(val[it], g["lambda_val"][it])


@pytest.mark.gpu
def test_gpu_shortchar_origin_ranges_and_other_approximations():
    """Origin sub-ranges give the rows of the full run; OneLine and None (sum over all lines) against the oracle."""
    from magritte_b200 import cabi
    from oracle import oracle as orc
    flat, _ = util.load_golden("wind3d_spiral_2spec")
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        for approx in (0, 1, 2):
            o = orc.Oracle(flat, approx=approx, store_u=False)
            o.shortchar()
            ctx.run_shortchar(approx)
            J = ctx.get_J()
            val, cnt = ctx.get_lambda()
            assert np.array_equal(J, o.J), approx
            assert np.array_equal(val, o.L_val[:, :, 0]) and np.array_equal(cnt, o.L_cnt), approx
        N = int(flat["npoints"])
        ctx.run_shortchar(2, (N // 3, N // 2))
        Jp = ctx.get_J()
        assert Jp.shape[0] == N // 2 - N // 3 and np.array_equal(Jp, J[N // 3:N // 2])
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"])
@pytest.mark.parametrize("device_lp", [True, False])
def test_gpu_host_model_shortchar_radiation_field_and_one_iteration(name, device_lp):
    """Model::compute_radiation_field_shortchar_order_0 and ONE iteration of compute_level_populations_shortchar through the C++ host
    Model: J bit for bit; populations (populations criterion) and Jeff of the interior points against the compiled reference (a
    point's statistical equilibrium reads its own Lambda element only, and the boundary points' elements are undefined there)."""
    from test_gpu_levelpop import prepared
    from test_host_model import assert_populations_close
    g = load_shortchar(name)
    m, flat, ref, lp = prepared(name)
    m.device_level_populations = device_lp
    m.compute_radiation_field_shortchar_order_0()
    assert np.array_equal(m.get_J(), g["J"])
    m.mutable_parameters().use_adaptive_Ng_acceleration = False
    assert m.compute_level_populations_shortchar(False, 1) == 1
    it = g["interior"]
    for l in range(len(lp["species"])):
        got = m.get_population(l)
        want = g[f"iter1_population_{l}"].reshape(got.shape)
        assert_populations_close(got[it], want[it], np.asarray(m.get_population_tot(l))[it], name)
        Jeff = np.asarray(m.get_Jeff(l))
        assert util.relerr(Jeff[it], g[f"iter1_Jeff_{l}"].reshape(Jeff.shape)[it]) <= 1e-12


@pytest.mark.gpu
def test_gpu_facade_shortchar():
    """The magritte.core surface: compute_radiation_field_shortchar_order_0 fills radiation.J and the Lambda lists."""
    import magritte_b200.core_compat as core
    from magritte_b200 import modelbuild as mb
    name = "velgrad_1d"
    g = load_shortchar(name)
    m = mb.build(core, core.IoPython("mem", "sc_facade"), "sc_facade", util.small_specs()[name])
    mb.prepare(m)
    assert m.compute_radiation_field_shortchar_order_0() == 0
    assert np.array_equal(np.array(m.radiation.J), g["J"])
    it = np.flatnonzero(g["interior"])
    Ls = m.lines.lineProducingSpecies[0].Lambda.Ls
    nr = m.lines.lineProducingSpecies[0].Lambda.nr
    for p in it[:50]:
        for k in range(g["lambda_val"].shape[1]):
            assert nr[p][k] == [p] and Ls[p][k] == [g["lambda_val"][p, k]]
