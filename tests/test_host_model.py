"""Host layer (magritte_b200/host/mgb_host.hpp through the pybind11 module magritte_b200.core) against the compiled reference's
golden vectors, WITHOUT a GPU: line widths, spectral discretisation, LTE populations, emissivity/opacity (bit-exact), and
compute_Jeff + the statistical-equilibrium step fed with the reference's own J and Lambda (tests/golden/levelpop_*.pkl)."""
import os
import pickle

import numpy as np
import pytest

import util

CASES = ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"]


def assert_populations_close(got, want, tot, ctx="", abs_tot=1e-15):
    """1e-10 relative (BASELINE.json north_star) plus an absolute term of 1e-15 of the point's total population: the two
    long-double LU factorisations (dense per block here, global SparseLU/COLAMD in the reference) agree to ~1e-17 of the
    total, which for a level holding less than 1e-7 of the molecules is more than 1e-10 of that level -- solver noise
    that the reference's result carries itself (it also ignores levels below 1e-10 of the total when testing convergence)."""
    want = np.asarray(want).reshape(got.shape)
    bound = 1e-10 * np.abs(want) + abs_tot * np.asarray(tot)[:, None]
    bad = np.abs(got - want) > bound
    assert not bad.any(), f"{ctx}: {bad.sum()} populations differ, worst {np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300))}"


def load_levelpop(name):
    with open(os.path.join(util.GOLDEN, f"levelpop_{name}.pkl"), "rb") as f:
        return pickle.load(f)


def host_model(name):
    from magritte_b200 import model as hm
    flat, ref = util.load_golden(name)
    lp = load_levelpop(name)
    m = hm.build(flat, lp["species"], lp["abundance"], lp["parameters"])
    return m, flat, ref, lp


@pytest.mark.parametrize("name", CASES)
def test_setup_matches_reference_bit_for_bit(name):
    m, flat, ref, lp = host_model(name)
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    assert np.array_equal(m.get_nu(), flat["nu"])
    assert np.array_equal(m.get_inverse_width(), flat["inverse_width"])
    for l, want in enumerate(lp["lte_population"]):
        assert np.array_equal(m.get_population(l).ravel(), want.ravel())
        assert np.array_equal(m.get_population_tot(l), flat["population_tot"][l])
    assert np.array_equal(m.get_emissivity(), lp["lte_emissivity"]) and np.array_equal(m.get_emissivity(), flat["emissivity"])
    assert np.array_equal(m.get_opacity(), lp["lte_opacity"]) and np.array_equal(m.get_opacity(), flat["opacity"])


@pytest.mark.parametrize("name", CASES)
def test_stateq_step_on_reference_radiation_field(name):
    """compute_Jeff (bit-exact) and one statistical-equilibrium update (block-wise dense LU vs the reference's global SparseLU,
    both in long double) on the J and Lambda of the reference's first iteration."""
    m, flat, ref, lp = host_model(name)
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    m.set_radiation_field(ref["ref_J"], ref["ref_lambda_val"], ref["ref_lambda_cnt"])
    m.compute_Jeff()
    for l, want in enumerate(lp["iter1_Jeff"]):
        assert np.array_equal(m.get_Jeff(l), want)
    m.compute_level_populations_from_stateq()
    for l, want in enumerate(lp["iter1_population"]):
        assert_populations_close(m.get_population(l), want, m.get_population_tot(l), name)
    nl = len(lp["iter1_population"])
    for l in range(nl):
        fnc, mean, mx = m.convergence(l)
        assert abs(mx - lp["iter1_error_max"][l]) <= 1e-9 * max(1.0, abs(mx))
        assert abs(mean - lp["iter1_error_mean"][l]) <= 1e-9 * max(1.0, abs(mean))


BANDED = ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"]


@pytest.mark.parametrize("name,solver", [(n, s) for n in BANDED for s in ("direct", "krylov", None)
                                         if not (s == "direct" and n.startswith("wind3d"))])
def test_banded_stateq_step_on_reference_radiation_field(name, solver, monkeypatch):
    """n_off_diag > 0: compute_Jeff over the whole Lambda lists (bit-exact) and the GLOBAL statistical-equilibrium system (blocks +
    off-block Lambda elements; Eigen::SparseLU in the reference) on the J and the lists of the reference's first iteration
    (tests/golden/offdiag_*.npz, offdiag_levelpop_*.npz). Both host solvers: the direct band LU (what 1D models get) and the
    block-Jacobi preconditioned GMRES refinement (what 3D point clouds get); solver = None is the automatic choice."""
    from magritte_b200 import model as hm
    if solver:
        monkeypatch.setenv("MGBH_STATEQ_SOLVER", solver)
    else:
        monkeypatch.delenv("MGBH_STATEQ_SOLVER", raising=False)
    flat, lists = util.load_golden_offdiag(name)
    lp = load_levelpop(name)
    gold = np.load(os.path.join(util.GOLDEN, f"offdiag_levelpop_{name}.npz"))
    prm = dict(lp["parameters"], n_off_diag=int(gold["n_off_diag"]), use_adaptive_Ng_acceleration=False)
    m = hm.build(flat, lp["species"], lp["abundance"], prm)
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    m.set_radiation_field_lists(lists["J"], lists["lambda_val"], lists["lambda_nr"], lists["lambda_cnt"])
    m.compute_Jeff()
    ns = int(gold["nspecies"])
    for l in range(ns):
        assert np.array_equal(m.get_Jeff(l), gold[f"iter1_Jeff_{l}"]), f"{name}: Jeff of species {l}"
    m.compute_level_populations_from_stateq()
    coupled = False
    for l in range(ns):
        assert_populations_close(m.get_population(l), gold[f"iter1_population_{l}"], m.get_population_tot(l), name)
        its, res = m.stateq_solver_info(l)
        assert res <= 1e-17, (name, l, its, res)  # component-wise backward error of the long-double solve
        coupled |= its > 0
        fnc, mean, mx = m.convergence(l)
        assert abs(mx - gold["iter1_error_max"][l]) <= 1e-9 * max(1.0, abs(mx))
        assert abs(mean - gold["iter1_error_mean"][l]) <= 1e-9 * max(1.0, abs(mean))
    if solver == "krylov" and name != "c1_all_constant":
        assert coupled, "the off-diagonal elements never entered the solve"
    if solver is None:
        assert coupled == name.startswith("wind3d"), "1D models take the direct band LU, 3D clouds the Krylov refinement"


@pytest.mark.parametrize("name", BANDED)
def test_banded_level_population_loop_driven_by_the_oracle(name):
    """The whole n_off_diag > 0 loop without a GPU: the radiation field and its banded Lambda lists from the CPU oracle, Jeff and the
    global statistical-equilibrium solve from the host layer, K iterations against the compiled reference's own loop
    (offdiag_levelpop_*.npz: final_population). Bound: 1e-10 relative + 1e-13 of the point's total -- the banded system is worse
    conditioned than the block-diagonal one (a point collects the Lambda elements of every list that names it), and the 1e-13
    differences between two correct radiation fields reach the small levels amplified; the step on the reference's OWN radiation
    field (the test above) holds the 1e-15 bound."""
    from magritte_b200 import model as hm
    from oracle import oracle as orc
    flat, _ = util.load_golden_offdiag(name)
    lp = load_levelpop(name)
    gold = np.load(os.path.join(util.GOLDEN, f"offdiag_levelpop_{name}.npz"))
    K = int(gold["n_off_diag"])
    m = hm.build(flat, lp["species"], lp["abundance"], dict(lp["parameters"], n_off_diag=K, use_adaptive_Ng_acceleration=False))
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    f2 = dict(flat)
    for _ in range(int(gold["iterations"])):
        f2["emissivity"], f2["opacity"] = m.get_emissivity(), m.get_opacity()
        o = orc.Oracle(f2, n_off_diag=K, store_u=False, threads=min(8, os.cpu_count() or 1))
        o.solve()
        m.set_radiation_field_lists(o.J, o.L_val, o.L_nr, o.L_cnt)
        m.compute_Jeff()
        m.compute_level_populations_from_stateq()
    for l in range(int(gold["nspecies"])):
        assert_populations_close(m.get_population(l), gold[f"final_population_{l}"], m.get_population_tot(l), name, abs_tot=1e-13)
        assert m.stateq_solver_info(l)[1] <= 1e-17


def test_errors_mirror_the_reference():
    m, flat, ref, lp = host_model("c1_all_constant")
    with pytest.raises(RuntimeError, match="Spectral discretisation was not set for Lines"):
        m.compute_level_populations(False, 1)
    with pytest.raises(RuntimeError, match="does not point in a direction"):
        m.compute_image_new(0.0, 0.0, 0.0, 4, 4)
    with pytest.raises(RuntimeError, match="At least a single frequency"):
        m.compute_spectral_discretisation(1.0e11, 2.0e11, 0)
