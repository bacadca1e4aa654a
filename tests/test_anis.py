"""Solver::solve_feautrier_order_2_anis (reference solver.tpp:365-456; Model::compute_radiation_field_feautrier_order_2_anis,
model.cpp:435-470): lspec.J and the anisotropic radiation-field tensor J2_0 .. J2_2_Im per line, from the same solves as the Feautrier
order-2 field. Goldens from the compiled reference: tests/golden/make_golden_anis.py.
CPU: the oracle's accumulation reproduces the reference bit for bit. GPU: the CUDA path <= 1e-10 against the reference (absolute floor
for the tensor elements, which cancel over the sphere) and bit-exact against the oracle on the two-sided solves."""
import os

import numpy as np
import pytest

import util

CASES = util.GOLDEN_CASES + ["wind3d_adaptive"]
Q = ("J", "J2_0", "J2_1_Re", "J2_1_Im", "J2_2_Re", "J2_2_Im")


def golden(name):
    z = np.load(os.path.join(util.GOLDEN, f"anis_{name}.npz"))
    return np.stack([z[q] for q in Q])


@pytest.mark.parametrize("name", CASES)
def test_oracle_anis_bit_exact(name):
    from oracle import oracle as orc
    flat, ref = util.load_golden(name)
    o = orc.Oracle(flat, store_u=True)
    o.solve()
    assert np.array_equal(o.anis(), golden(name))


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_gpu_anis(name):
    from magritte_b200 import cabi
    from magritte_b200.flatten import approx_of
    from oracle import oracle as orc
    flat, ref = util.load_golden(name)
    want = golden(name)
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        ctx.set_params(anisotropic=1)
        ctx.run(approx_of(flat["params"]))
        got = ctx.get_J_anis()
        ctx.set_params(anisotropic=0)
        ctx.run(approx_of(flat["params"]))
        with pytest.raises(cabi.MgbError, match="anisotropic"):
            ctx.get_J_anis()
    finally:
        ctx.close()
    assert util.relerr(got[0], want[0]) <= 1e-10
    scale = np.abs(want[0])  # the tensor elements are differences of O(J) terms over the sphere
    assert np.all(np.abs(got[1:] - want[1:]) <= 1e-10 * np.abs(want[1:]) + 1e-12 * scale[None])
    o = orc.Oracle(flat, store_u=True)
    o.solve(twosided=True)
    assert np.array_equal(got, o.anis())


@pytest.mark.gpu
def test_anis_through_host_model_and_facade():
    from magritte_b200 import core_compat as core, modelbuild as mb
    name = "wind3d_small"
    want = golden(name)
    m = mb.build(core, core.IoPython("mem", "anis_t"), "anis_t", util.small_specs()[name])
    mb.prepare(m)
    assert m.compute_radiation_field_feautrier_order_2_anis() == 0
    ls = m.lines.lineProducingSpecies[0]
    assert util.relerr(np.array(ls.J), want[0]) <= 1e-10
    for i, q in enumerate(Q[1:], start=1):
        assert np.all(np.abs(np.array(getattr(ls, q)) - want[i]) <= 1e-10 * np.abs(want[i]) + 1e-12 * np.abs(want[0]))
