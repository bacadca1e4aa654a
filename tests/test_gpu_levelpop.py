"""Model::compute_level_populations through the host layer (magritte_b200.core.Model -> C ABI -> CUDA) against the compiled
reference (tests/golden/levelpop_*.pkl): K iterations without Ng acceleration at 1e-10, and the default adaptive-Ng run
(same iteration count, populations at a looser bound because the Ng coefficients come from an ill-conditioned Gram matrix)."""
import numpy as np
import pytest

import util
from test_host_model import CASES, assert_populations_close, host_model

pytestmark = pytest.mark.gpu


def prepared(name):
    m, flat, ref, lp = host_model(name)
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    return m, flat, ref, lp


@pytest.mark.parametrize("name", CASES)
def test_K_iterations_without_Ng(name):
    m, flat, ref, lp = prepared(name)
    m.device_level_populations = False  # the host twin (long double); the device path: tests/test_gpu_levelpop_device.py
    m.mutable_parameters().use_adaptive_Ng_acceleration = False
    it = m.compute_level_populations(False, lp["K"])
    assert it == lp["K"]
    for l, want in enumerate(lp["noNg_population"]):
        assert_populations_close(m.get_population(l), want, m.get_population_tot(l), name)
        assert util.relerr(m.get_Jeff(l), lp["noNg_Jeff"][l]) <= 1e-10
    assert util.relerr(m.get_J()[::7], lp["noNg_J"]) <= 1e-10
    assert np.allclose(m.error_max, lp["noNg_error_max"], rtol=1e-8, atol=1e-14)
    assert m.seconds_radiation_field > 0


@pytest.mark.parametrize("name", ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small"])
def test_default_adaptive_Ng_run(name):
    m, flat, ref, lp = prepared(name)
    m.device_level_populations = False
    it = m.compute_level_populations(True, 60)
    assert it == lp["Ng_iterations"]
    for l, want in enumerate(lp["Ng_population"]):
        got = m.get_population(l)
        want = want.reshape(got.shape)
        tot = m.get_population_tot(l)[:, None]
        assert np.all(np.abs(got - want) <= 1e-6 * np.abs(want) + 1e-12 * tot), name


@pytest.mark.parametrize("name", ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"])
def test_banded_lambda_iterations(name):
    """parameters.n_off_diag > 0 through the host Model: the radiation field's banded Lambda lists come from the GPU
    (mgb_get_lambda_lists), Jeff and the GLOBAL statistical-equilibrium system (neighbouring points couple) are solved on the host
    (block-Jacobi preconditioned GMRES in long double); K iterations against the compiled reference's SparseLU loop
    (tests/golden/offdiag_levelpop_*.npz)."""
    import os
    m, flat, ref, lp = prepared(name)
    gold = np.load(os.path.join(util.GOLDEN, f"offdiag_levelpop_{name}.npz"))
    P = m.mutable_parameters()
    P.use_adaptive_Ng_acceleration = False
    P.n_off_diag = int(gold["n_off_diag"])
    K = int(gold["iterations"])
    it = m.compute_level_populations(False, K)
    assert it == K
    for l in range(int(gold["nspecies"])):
        # 1e-13 of the point's total: see test_banded_level_population_loop_driven_by_the_oracle (tests/test_host_model.py)
        assert_populations_close(m.get_population(l), gold[f"final_population_{l}"], m.get_population_tot(l), name, abs_tot=1e-13)
        assert m.stateq_solver_info(l)[1] <= 1e-17
    assert np.allclose(m.error_max, gold["final_error_max"], rtol=1e-8, atol=1e-14)


def test_image_through_host_model():
    m, flat, ref, lp = prepared("wind3d_small")
    fl, img = util.load_golden_image("wind3d_small")
    nu = fl["nu_image"]
    m.compute_spectral_discretisation(float(nu[0]), float(nu[-1]), len(nu))
    d = img["directions"][1]
    m.compute_image_new(float(d[0]), float(d[1]), float(d[2]), img["nx"], img["ny"])
    X, Y, I = img["images"][1]
    im = m.images[-1]
    assert np.array_equal(im.ImX, X) and np.array_equal(im.ImY, Y)
    assert util.relerr(im.I, I) <= 1e-10
    # the optical-depth image of the same pixels through the same host method family (Model::compute_image_optical_depth_new)
    m.compute_image_optical_depth_new(float(d[0]), float(d[1]), float(d[2]), img["nx"], img["ny"])
    tau = m.images[-1]
    assert tau.optical_depth and np.array_equal(np.asarray(tau.I).reshape(-1), util.load_golden_image_tau("wind3d_small")[1][2].reshape(-1))
