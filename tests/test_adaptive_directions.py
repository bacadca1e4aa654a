"""Rays::use_adaptive_directions -- a ray set (directions and weights) per point: reference src/model/geometry/rays/rays.hpp:23-31,
rays.cpp:37-100; Geometry::get_next<true> / get_shift<..., true> (geometry.tpp:270-277, 428-438); Rays::get_weight<true> in the J and
Lambda accumulation (solver.tpp:504-506, 1777); selected at model.cpp:357-389. Golden vectors from the compiled reference:
tests/golden/make_golden_adaptive.py.

CPU: the oracle reproduces the reference bit for bit. GPU (-m gpu): J, Lambda and column densities against the reference, bit-exact
against the oracle's two-sided restatement, every ray path against the oracle, and K level-population iterations through the host Model
and through the magritte.core-compatible facade (modelbuild.build() unchanged)."""
import os
import pickle

import numpy as np
import pytest

import util
from test_host_model import assert_populations_close

NAME = "wind3d_adaptive"


def load_lp():
    with open(os.path.join(util.GOLDEN, f"levelpop_{NAME}.pkl"), "rb") as f:
        return pickle.load(f)


def test_golden_is_adaptive_and_oracle_bit_exact():
    from oracle import oracle as orc
    flat, ref = util.load_golden(NAME)
    N, R = int(flat["npoints"]), int(flat["nrays"])
    assert flat["use_adaptive_directions"] and flat["direction"].shape == (N * R, 3) and flat["weight"].shape == (N * R,)
    assert not np.array_equal(flat["direction"][:R], flat["direction"][R:2 * R])  # the sets really differ from point to point
    o = orc.Oracle(flat, store_u=False)
    J = o.solve()
    assert np.array_equal(J, ref["ref_J"])
    assert np.array_equal(o.L_cnt, ref["ref_lambda_cnt"]) and np.array_equal(o.L_val[:, :, 0], ref["ref_lambda_val"])
    assert np.array_equal(o.set_column(ref["ref_density"]), ref["ref_column"])


def test_synth_flat_of_adaptive_spec():
    from magritte_b200 import modelbuild as mb, synth
    flat, ref = util.load_golden(NAME)
    mine = synth.flat_from_spec(mb.with_adaptive_directions(util.small_specs()["wind3d_small"]))
    for k in ("direction", "weight", "antipod", "nu", "emissivity", "opacity"):
        assert np.array_equal(mine[k], flat[k]), k
    assert mine["nrays"] == flat["nrays"] and mine["use_adaptive_directions"]


@pytest.mark.gpu
def test_J_lambda_column_against_reference_and_twosided_oracle():
    from magritte_b200 import cabi
    from magritte_b200.flatten import approx_of
    from oracle import oracle as orc
    flat, ref = util.load_golden(NAME)
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        st = ctx.run(approx_of(flat["params"]))
        J = ctx.get_J()
        val, cnt = ctx.get_lambda()
        col = ctx.set_column(ref["ref_density"])
    finally:
        ctx.close()
    assert util.relerr(J, ref["ref_J"]) <= 1e-10
    assert np.array_equal(cnt, ref["ref_lambda_cnt"]) and util.relerr(val, ref["ref_lambda_val"]) <= 1e-10
    assert np.array_equal(col, ref["ref_column"])
    o = orc.Oracle(flat, store_u=False)
    assert np.array_equal(J, o.solve(twosided=True))
    assert np.array_equal(val, o.L_val[:, :, 0])
    # the same cloud WITHOUT adaptive directions gives another J: the per-point sets are really used
    base, _ = util.load_golden("wind3d_small")
    assert util.relerr(J, _["ref_J"]) > 1e-3


@pytest.mark.gpu
def test_every_ray_path_against_oracle():
    from magritte_b200 import cabi
    from oracle import oracle as orc
    flat, ref = util.load_golden(NAME)
    o = orc.Oracle(flat, store_u=False)
    cap = max(8, o.max_length())
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        for org in range(0, int(flat["npoints"]), 3):
            for rr in range(int(flat["hnrays"])):
                for l in (0, int(flat["nlines"]) - 1):
                    util.assert_trace_equal(ctx.trace(org, rr, l, cap=2 * cap + 1), o.trace(org, rr, l, cap=cap), f"({org},{rr},{l})")
    finally:
        ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("device_lp", [True, False])
def test_level_populations_through_host_model(device_lp):
    from magritte_b200 import model as hm
    flat, ref = util.load_golden(NAME)
    lp = load_lp()
    m = hm.build(flat, lp["species"], lp["abundance"], lp["parameters"])
    m.device_level_populations = device_lp
    m.compute_spectral_discretisation(); m.compute_inverse_line_widths(); m.compute_LTE_level_populations()
    m.mutable_parameters().use_adaptive_Ng_acceleration = False
    assert m.compute_level_populations(False, lp["K"]) == lp["K"]
    for l, want in enumerate(lp["noNg_population"]):
        assert_populations_close(m.get_population(l), want, m.get_population_tot(l), NAME)
        assert util.relerr(m.get_Jeff(l), lp["noNg_Jeff"][l]) <= 1e-10
    assert util.relerr(m.get_J()[::7], lp["noNg_J"]) <= 1e-10
    assert np.allclose(m.error_max, lp["noNg_error_max"], rtol=1e-8, atol=1e-14)


@pytest.mark.gpu
def test_same_script_through_the_compat_facade():
    from magritte_b200 import core_compat as core, modelbuild as mb
    flat, ref = util.load_golden(NAME)
    spec = mb.with_adaptive_directions(util.small_specs()["wind3d_small"])
    m = mb.build(core, core.IoPython("mem", NAME), NAME, spec)
    assert m.geometry.rays.use_adaptive_directions and m.parameters.nrays() == int(flat["nrays"])
    mb.prepare(m)
    assert m.compute_radiation_field_feautrier_order_2() == 0
    assert util.relerr(np.array(m.radiation.J), ref["ref_J"]) <= 1e-10
