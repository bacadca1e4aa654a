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
        val, cnt = ctx.get_lambda_diag()
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
        for approx in (cabi.APPROX_NONE, cabi.APPROX_ONELINE, cabi.APPROX_CLOSELINES):
            o = orc.Oracle(flat, approx=approx, store_u=False)
            o.shortchar()
            ctx.run_shortchar(approx)
            J = ctx.get_J()
            val, cnt = ctx.get_lambda_diag()
            assert np.array_equal(J, o.J), approx
            assert np.array_equal(val, o.L_val[:, :, 0]) and np.array_equal(cnt, o.L_cnt), approx
        N = int(flat["npoints"])
        ctx.run_shortchar(cabi.APPROX_CLOSELINES, (N // 3, N // 2))
        Jp = ctx.get_J()
        assert np.array_equal(Jp[N // 3:N // 2], J[N // 3:N // 2])
    finally:
        ctx.close()
