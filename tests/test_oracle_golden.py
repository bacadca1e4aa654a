"""The oracle (oracle/feautrier_oracle.c, a CPU restatement of the reference's path) against the golden
vectors generated from the compiled, unmodified reference (tests/golden/make_golden.py): bit for bit."""
import numpy as np
import pytest

import util
from oracle import oracle as orc


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_oracle_bit_exact(name):
    flat, ref = util.load_golden(name)
    o = orc.Oracle(flat, store_u=True)
    assert np.array_equal(o.dshift_max, ref["ref_dshift_max"])
    assert 2 * o.max_length() + 1 == int(ref["ref_length"])  # Solver::setup: length = 2*max+1 (solver.tpp:4)
    J = o.solve()
    assert np.array_equal(J, ref["ref_J"]), f"max rel err {util.relerr(J, ref['ref_J'])}"
    if "ref_u" in ref:
        assert np.array_equal(o.u, ref["ref_u"])
    assert np.array_equal(o.L_cnt, ref["ref_lambda_cnt"])
    assert np.array_equal(o.L_val[:, :, 0], ref["ref_lambda_val"])
    for sel, want in util.golden_traces(ref):
        util.assert_trace_equal(o.trace(*sel), want, f"{name} {sel}")


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_twosided_reordering_error(name):
    """The product eliminates the tridiagonal system from both ends (mgb_solve.cuh). Restated on the CPU, that
    re-ordering must stay far inside the 1e-10 tolerance of BASELINE.json's north_star against the reference."""
    flat, ref = util.load_golden(name)
    o = orc.Oracle(flat, store_u=True)
    J = o.solve(twosided=True).copy()
    assert util.relerr(J, ref["ref_J"]) < 1e-12
    assert util.relerr(o.L_val[:, :, 0], ref["ref_lambda_val"]) < 1e-12
    if "ref_u" in ref:
        assert util.relerr(o.u, ref["ref_u"]) < 1e-12


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_oracle_image_bit_exact(name):
    """The oracle's restatement of the new imager against Image.ImX / ImY / I of the compiled reference: bit for bit."""
    flat, img = util.load_golden_image(name)
    o = orc.Oracle(flat, store_u=False)
    for d, (X, Y, I) in zip(img["directions"], img["images"]):
        r = o.image(d, img["nx"], img["ny"], img["dimension"])
        assert np.array_equal(r["ImX"], X) and np.array_equal(r["ImY"], Y)
        assert np.array_equal(r["I"], I), f"{name} {d}: max rel err {util.relerr(r['I'], I)}"


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_oracle_optical_depth_image_bit_exact(name):
    """Solver::image_optical_depth_new_imager restated: Image.I (optical depth) of the compiled reference, bit for bit."""
    flat, img = util.load_golden_image(name)
    o = orc.Oracle(flat, store_u=False)
    for d, (X, Y, T) in zip(img["directions"], util.load_golden_image_tau(name)):
        r = o.image(d, img["nx"], img["ny"], img["dimension"], optical_depth=True)
        assert np.array_equal(r["ImX"], X) and np.array_equal(r["ImY"], Y)
        assert np.array_equal(r["I"], T), f"{name} {d}: max rel err {util.relerr(r['I'], T)}"


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_oracle_banded_lambda_bit_exact(name):
    """n_off_diag > 0: the oracle's Lambda lists equal the compiled reference's, element for element in insertion order."""
    flat, ref = util.load_golden_offdiag(name)
    o = orc.Oracle(flat, store_u=False)
    J = o.solve()
    assert np.array_equal(J, ref["J"])
    cap = ref["lambda_nr"].shape[2]
    util.lambda_lists_close(o.L_val[:, :, :cap], o.L_nr[:, :, :cap], o.L_cnt, ref, 0)


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_oracle_column_bit_exact(name):
    """orc_set_column against Model::column of the compiled reference (tests/golden/make_golden_column.py): bit for bit."""
    flat, _ = util.load_golden(name)
    density, column = util.load_golden_column(name)
    o = orc.Oracle(flat, store_u=False)
    assert np.array_equal(o.set_column(density), column)
    sub = o.set_column(density, o_range=(3, 17))
    assert np.array_equal(sub, column[:, 3:17])
