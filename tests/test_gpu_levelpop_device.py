"""The level-population half of an ALI iteration ON THE DEVICE (mgb_lp_*, magritte_b200/csrc/mgb_levelpop.cu) against the compiled
reference's golden vectors (tests/golden/levelpop_*.pkl, sparse_*.pkl):

  * compute_Jeff + ONE statistical-equilibrium update fed with the reference's own J and Lambda, through the C ABI: Jeff at 1e-15
    relative, populations at the populations criterion;
  * WHY that criterion has an absolute term: for sampled points the same rate matrix is solved in 50-digit arithmetic (mpmath) and
    the device result (double-double elimination) must be CLOSER to it than the reference's long-double SparseLU result is -- the
    residual disagreement is the reference's own rounding, not this solver's;
  * the whole loop through the host Model (device_level_populations = True): K iterations without Ng acceleration, the adaptive-Ng
    run, the memory-sparse variants;
  * that an iteration moves no per-point table across PCIe (the line state the solver reads is the one the device wrote)."""
import numpy as np
import pytest

import util
from test_host_model import CASES, assert_populations_close, load_levelpop
from test_gpu_levelpop import prepared
from test_gpu_sparse import load_sparse

pytestmark = pytest.mark.gpu


def device_ctx(name):
    from magritte_b200 import cabi
    flat, ref = util.load_golden(name)
    lp = load_levelpop(name)
    ctx = cabi.Context(0)
    ctx.set_model(flat)
    ctx.lp_setup(lp["species"], lp["abundance"], lp["parameters"])
    for l, pop in enumerate(lp["lte_population"]):
        ctx.lp_set_populations(l, np.asarray(pop, dtype=np.float64).reshape(int(flat["npoints"]), -1), flat["population_tot"][l])
    return ctx, flat, ref, lp


@pytest.mark.parametrize("name", CASES)
def test_line_state_from_device_populations(name):
    """Lines::set_emissivity_and_opacity / set_inverse_width on the device from the LTE populations: inverse widths bit for bit,
    emissivity within one rounding (the reference rounds a long double product to double, the device a double-double one); the
    opacity is a difference of two such products and the golden populations are the reference's long doubles rounded to double, so the
    cancellation (h nu / k T of the transition) amplifies that rounding."""
    ctx, flat, ref, lp = device_ctx(name)
    try:
        ctx.lp_refresh_line_state()
        e, o, w = ctx.lp_get_line_state()
    finally:
        ctx.close()
    assert np.array_equal(w, flat["inverse_width"])
    assert util.relerr(e, flat["emissivity"]) <= 2.3e-16 and util.relerr(o, flat["opacity"]) <= 1e-12


@pytest.mark.parametrize("name", CASES)
def test_stateq_step_on_reference_radiation_field(name):
    ctx, flat, ref, lp = device_ctx(name)
    try:
        ctx.set_radiation_field(ref["ref_J"], ref["ref_lambda_val"], ref["ref_lambda_cnt"])
        ctx.lp_compute_Jeff()
        conv, ms = ctx.lp_update_statistical_equilibrium(pop_prec=float(lp["parameters"].get("pop_prec", 1e-6)))
        for l, want in enumerate(lp["iter1_Jeff"]):
            Jlin, Jeff = ctx.lp_get_Jeff(l)
            assert util.relerr(Jeff, want) <= 1e-14
        for l, want in enumerate(lp["iter1_population"]):
            hi, lo = ctx.lp_get_populations(l)
            assert_populations_close(hi, want, flat["population_tot"][l], name)
            fnc, mean, mx = conv[l]
            assert abs(mx - lp["iter1_error_max"][l]) <= 1e-9 * max(1.0, abs(mx))
            assert abs(mean - lp["iter1_error_mean"][l]) <= 1e-9 * max(1.0, abs(mean))
    finally:
        ctx.close()


def _expf(x):
    """glibc's float expf (what collisionPartner.tpp:36 calls), not numpy's."""
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    libm.expf.restype, libm.expf.argtypes = ctypes.c_float, [ctypes.c_float]
    return libm.expf(ctypes.c_float(x))


def exact_populations(flat, lp, ref, l, points):
    """The rate equations of `points` for species l, entries formed in double exactly like the reference (lineProducingSpecies.tpp:
    410-500), solved in 50-digit arithmetic."""
    import mpmath as mp
    mp.mp.dps = 50
    d = lp["species"][l]
    nlev, nrad = int(d["nlev"]), int(d["nrad"])
    irad, jrad = np.asarray(d["irad"]), np.asarray(d["jrad"])
    A, Ba, Bs, freq = (np.asarray(d[k], dtype=np.float64) for k in ("A", "Ba", "Bs", "frequency"))
    l0 = int(flat["nrad_cum"][l]) if "nrad_cum" in flat else int(np.concatenate([[0], np.cumsum(flat["nrad"])])[l])
    Jeff = np.asarray(lp["iter1_Jeff"][l]).reshape(int(flat["npoints"]), nrad)
    T, ab = np.asarray(flat["temperature"]), np.asarray(lp["abundance"])
    out = {}
    for p in points:
        M = mp.zeros(nlev, nlev)
        y = mp.zeros(nlev, 1)

        def add(r, c, v):
            M[r, c] += mp.mpf(float(v))
        for k in range(nrad):
            v_IJ = np.float64(A[k]) + np.float64(Bs[k]) * np.float64(Jeff[p, k])
            v_JI = np.float64(Ba[k]) * np.float64(Jeff[p, k])
            I, J = int(irad[k]), int(jrad[k])
            if J != nlev - 1:
                add(J, I, +v_IJ); add(J, J, -v_JI)
            if I != nlev - 1:
                add(I, J, +v_JI); add(I, I, -v_IJ)
        for k in range(nrad):
            for _ in range(int(ref["ref_lambda_cnt"][p, l0 + k])):
                v_IJ = -np.float64(ref["ref_lambda_val"][p, l0 + k]) * np.float64(flat["opacity"][p, l0 + k]) / np.float64(freq[k])
                I, J = int(irad[k]), int(jrad[k])
                if J != nlev - 1:
                    add(J, I, +v_IJ)
                if I != nlev - 1:
                    add(I, I, -v_IJ)
        for cp in d["colpar"]:
            abn = np.float64(ab[p, int(cp["num_col_partner"])])
            tmp = np.asarray(cp["tmp"], dtype=np.float64)
            if str(cp["orth_or_para_H2"]) != "n":
                frac = np.float64(1.0) / (np.float64(1.0) + np.float64(9.0) * np.float64(_expf(-170.5 / T[p])))
                abn = abn * (np.float64(1.0) - frac) if str(cp["orth_or_para_H2"]) == "o" else abn * frac
            Ce, Cd = np.asarray(cp["Ce"], dtype=np.float64), np.asarray(cp["Cd"], dtype=np.float64)
            if T[p] >= tmp[-1]:
                ce, cd = Ce[-1], Cd[-1]
            elif T[p] <= tmp[0]:
                ce, cd = Ce[0], Cd[0]
            else:
                t = int(np.searchsorted(tmp, T[p], side="left"))  # first tabulated temperature >= T
                step = (T[p] - tmp[t - 1]) / (tmp[t] - tmp[t - 1])
                ce = Ce[t - 1] + (Ce[t] - Ce[t - 1]) * step
                cd = Cd[t - 1] + (Cd[t] - Cd[t - 1]) * step
            for k in range(len(cp["icol"])):
                v_IJ, v_JI = cd[k] * abn, ce[k] * abn
                I, J = int(cp["icol"][k]), int(cp["jcol"][k])
                if J != nlev - 1:
                    add(J, I, +v_IJ); add(J, J, -v_JI)
                if I != nlev - 1:
                    add(I, J, +v_JI); add(I, I, -v_IJ)
        for i in range(nlev):
            add(nlev - 1, i, 1.0)
        y[nlev - 1] = mp.mpf(float(flat["population_tot"][l][p]))
        x = mp.lu_solve(M, y)
        out[p] = [x[i] for i in range(nlev)]
    return out


@pytest.mark.parametrize("name", ["c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"])
def test_device_solution_is_closer_to_exact_than_the_reference(name):
    """The absolute term of assert_populations_close is the REFERENCE's rounding: against the 50-digit solution of the same matrix
    the device populations must be at least as close as the reference's long-double SparseLU populations, level by level (up to the
    1e-16 relative rounding of reading the device's {hi, lo} pair and the golden's double)."""
    import mpmath as mp
    ctx, flat, ref, lp = device_ctx(name)
    try:
        ctx.set_radiation_field(ref["ref_J"], ref["ref_lambda_val"], ref["ref_lambda_cnt"])
        ctx.lp_compute_Jeff()
        ctx_Jeff = [ctx.lp_get_Jeff(l)[1] for l in range(len(lp["species"]))]
        # the negative-population clean-up renormalises; compare the raw solutions only where nothing was clamped
        ctx.lp_update_statistical_equilibrium()
        N = int(flat["npoints"])
        pts = sorted(set(np.linspace(0, N - 1, 7).astype(int).tolist()))
        worst_dev = worst_ref = 0.0
        for l in range(len(lp["species"])):
            assert np.array_equal(ctx_Jeff[l], np.asarray(lp["iter1_Jeff"][l]).reshape(ctx_Jeff[l].shape)) or util.relerr(ctx_Jeff[l], lp["iter1_Jeff"][l]) <= 1e-14
            hi, lo = ctx.lp_get_populations(l)
            want = np.asarray(lp["iter1_population"][l]).reshape(hi.shape)
            exact = exact_populations(flat, lp, ref, l, pts)
            for p in pts:
                if min(exact[p]) < 0:
                    continue  # clamped and renormalised by correct_negative_populations on both sides
                tot = float(flat["population_tot"][l][p])
                for i, xe in enumerate(exact[p]):
                    e_dev = abs(mp.mpf(float(hi[p, i])) + mp.mpf(float(lo[p, i])) - xe)
                    e_ref = abs(mp.mpf(float(want[p, i])) - xe)
                    # the golden is the reference's long double rounded to double (2^-53), and correct_negative_populations rescales every
                    # level by population_tot / (a sum rounded to double) on both sides (another 2^-52)
                    slack = mp.mpf(4.5e-16) * abs(xe) + mp.mpf(1e-30) * tot
                    assert e_dev <= e_ref + slack, (name, l, p, i, float(e_dev), float(e_ref), float(xe))
                    worst_dev = max(worst_dev, float(e_dev / tot)); worst_ref = max(worst_ref, float(e_ref / tot))
        print(f"{name}: worst |x - exact| / total  device {worst_dev:.2e}  reference (as double) {worst_ref:.2e}")
    finally:
        ctx.close()


@pytest.mark.parametrize("name", CASES)
def test_K_iterations_on_the_device(name):
    m, flat, ref, lp = prepared(name)
    assert m.device_level_populations
    m.mutable_parameters().use_adaptive_Ng_acceleration = False
    assert m.compute_level_populations(False, lp["K"]) == lp["K"]
    for l, want in enumerate(lp["noNg_population"]):
        assert_populations_close(m.get_population(l), want, m.get_population_tot(l), name)
        assert util.relerr(m.get_Jeff(l), lp["noNg_Jeff"][l]) <= 1e-10
    assert util.relerr(m.get_J()[::7], lp["noNg_J"]) <= 1e-10
    assert np.allclose(m.error_max, lp["noNg_error_max"], rtol=1e-8, atol=1e-14)


@pytest.mark.parametrize("name", CASES)
def test_device_and_host_loops_agree(name):
    """The host twin (long double, magritte_b200/host/mgb_host.hpp) and the device path run the same K iterations."""
    a, flat, ref, lp = prepared(name)
    b, _, _, _ = prepared(name)
    b.device_level_populations = False
    for m in (a, b):
        m.mutable_parameters().use_adaptive_Ng_acceleration = False
        m.compute_level_populations(False, lp["K"])
    for l in range(len(lp["noNg_population"])):
        assert_populations_close(a.get_population(l), b.get_population(l), a.get_population_tot(l), name)
        assert util.relerr(a.get_Jeff(l), b.get_Jeff(l)) <= 1e-12
    assert util.relerr(a.get_J(), b.get_J()) <= 1e-12
    assert util.relerr(a.get_emissivity(), b.get_emissivity()) <= 1e-10 and util.relerr(a.get_opacity(), b.get_opacity()) <= 1e-9


@pytest.mark.parametrize("name", ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small"])
def test_adaptive_Ng_on_the_device(name):
    m, flat, ref, lp = prepared(name)
    it = m.compute_level_populations(True, 60)
    assert it == lp["Ng_iterations"]
    for l, want in enumerate(lp["Ng_population"]):
        got = m.get_population(l)
        want = want.reshape(got.shape)
        tot = m.get_population_tot(l)[:, None]
        assert np.all(np.abs(got - want) <= 1e-6 * np.abs(want) + 1e-12 * tot), name


@pytest.mark.parametrize("name", ["c2_vanzadelhoff", "wind3d_small"])
def test_fixed_order_Ng_on_the_device(name):
    """use_adaptive_Ng_acceleration = False with Ng on: an Ng step of order mem_limit after mem_limit normal iterations
    (model.cpp:583-600). Device and host twin must take the same path and agree."""
    out = []
    for dev in (True, False):
        m, flat, ref, lp = prepared(name)
        m.device_level_populations = dev
        P = m.mutable_parameters()
        P.use_adaptive_Ng_acceleration = False
        P.Ng_acceleration_mem_limit = 3
        it = m.compute_level_populations(True, 5)
        out.append((it, [m.get_population(l) for l in range(len(lp["noNg_population"]))], m.get_population_tot(0), list(m.error_max)))
    assert out[0][0] == out[1][0]
    for a, b in zip(out[0][1], out[1][1]):
        assert np.all(np.abs(a - b) <= 1e-7 * np.abs(b) + 1e-13 * out[0][2][:, None]), name
    assert np.allclose(out[0][3], out[1][3], rtol=1e-6, atol=1e-12)


@pytest.mark.parametrize("name", CASES)
def test_K_sparse_iterations_on_the_device(name):
    m, flat, ref, lp = prepared(name)
    sp = load_sparse(name)
    m.mutable_parameters().use_adaptive_Ng_acceleration = False
    assert m.compute_level_populations_sparse(False, sp["K"]) == sp["K"]
    for l, want in enumerate(sp["noNg_population"]):
        assert_populations_close(m.get_population(l), want, m.get_population_tot(l), name, abs_tot=1e-13)  # the reference's QR noise, test_gpu_sparse.py
        assert util.relerr(m.get_Jeff(l), sp["noNg_Jeff"][l]) <= 1e-10
    assert np.allclose(m.error_max, sp["noNg_error_max"], rtol=1e-8, atol=1e-14)
