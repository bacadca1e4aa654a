"""Parity of the CUDA path (through the C ABI of include/magritte_b200.h) with the reference.

  * ray paths (nr, nr_interp, interp_factor, shift, dZ): bit-exact against the golden vectors dumped from the
    compiled reference and against the oracle;
  * J, Lambda, u: relative error <= 1e-10 (BASELINE.json north_star) against the reference's golden outputs, and
    BIT-EXACT against the oracle's restatement of the product's two-sided elimination order.
"""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

TOL = 1e-10  # north_star: "J, level populations and images within a relative tolerance of 1e-10 in fp64"


@pytest.fixture(scope="module")
def ctx():
    from magritte_b200 import cabi
    c = cabi.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_ray_paths_bit_exact(ctx, name):
    flat, ref = util.load_golden(name)
    ctx.set_model(flat)
    for sel, want in util.golden_traces(ref):
        util.assert_trace_equal(ctx.trace(*sel), want, f"{name} {sel}")


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
@pytest.mark.parametrize("cached", [False, True])
def test_J_lambda_against_reference(ctx, name, cached):
    from magritte_b200.flatten import approx_of
    flat, ref = util.load_golden(name)
    ctx.set_model(flat)
    ctx.set_params(store_intensities=1)
    approx = approx_of(flat["params"])
    st = ctx.run(approx)
    if cached:
        st = ctx.run(approx)
        assert st.paths_cached == 1
    J = ctx.get_J()
    assert util.relerr(J, ref["ref_J"]) <= TOL
    val, cnt = ctx.get_lambda()
    assert np.array_equal(cnt, ref["ref_lambda_cnt"])
    assert util.relerr(val, ref["ref_lambda_val"]) <= TOL
    if "ref_u" in ref:
        assert util.relerr(ctx.get_u(), ref["ref_u"]) <= TOL
    ctx.set_params(store_intensities=0)
    assert st.n_solves == flat["npoints"] * flat["hnrays"] * flat["nfreqs"]


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
@pytest.mark.parametrize("approx", [0, 1, 2])
def test_bit_exact_against_twosided_oracle(ctx, name, approx):
    """All three ApproximationTypes; the oracle's two-sided variant performs the kernel's operations in the kernel's order."""
    from oracle import oracle as orc
    flat, ref = util.load_golden(name)
    o = orc.Oracle(flat, approx=approx, store_u=False)
    Jo = o.solve(twosided=True).copy()
    ctx.set_model(flat)
    st = ctx.run(approx)
    J = ctx.get_J()
    val, cnt = ctx.get_lambda()
    assert st.n_steps == o.m.stat_steps and st.n_entries == o.m.stat_entries
    assert np.array_equal(J, Jo), f"max rel err {util.relerr(J, Jo)}"
    assert np.array_equal(cnt, o.L_cnt)
    assert np.array_equal(val, o.L_val[:, :, 0])
    # ... and the reference's own order stays within tolerance for every approximation
    Jr = o.solve(twosided=False)
    assert util.relerr(J, Jr) <= TOL


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_all_solve_kernel_generations_bit_identical(ctx, name, monkeypatch):
    """MGB_SOLVE_KERNEL = 1 (k_solve), 3 (k_solve3) and 4 (the streaming pipeline of mgb_solve4.cuh, the default): J, u, Lambda and the
    work counters must agree bit for bit, for all three ApproximationTypes; the pipeline also with a record budget so small that the
    origins are taken in many record ranges (and the range-halving path runs)."""
    flat, _ = util.load_golden(name)
    ctx.set_model(flat)
    ctx.set_params(store_intensities=1)
    try:
        for approx in (2, 1, 0):
            monkeypatch.setenv("MGB_SOLVE_KERNEL", "1")
            st1 = ctx.run(approx)
            J1, u1 = ctx.get_J().copy(), ctx.get_u().copy()
            l1, c1 = ctx.get_lambda()
            assert ctx.pipeline_info()["kernel"] == 1
            for kernel, budget in (("3", None), ("4", None), ("4", str(200 * 1024))):
                monkeypatch.setenv("MGB_SOLVE_KERNEL", kernel)
                if budget:
                    monkeypatch.setenv("MGB_REC_BYTES", budget)
                else:
                    monkeypatch.delenv("MGB_REC_BYTES", raising=False)
                st = ctx.run(approx)
                info = ctx.pipeline_info()
                assert info["kernel"] == int(kernel)
                if budget and flat["npoints"] > 100:
                    assert info["n_ranges"] > 3
                l, c = ctx.get_lambda()
                assert np.array_equal(ctx.get_J(), J1) and np.array_equal(ctx.get_u(), u1), (name, approx, kernel, budget)
                assert np.array_equal(l, l1) and np.array_equal(c, c1)
                assert st.n_entries == st1.n_entries and st.n_large_shift == st1.n_large_shift and st.n_solves == st1.n_solves
    finally:
        ctx.set_params(store_intensities=0)


def test_chunked_run_equals_single_chunk(ctx):
    """Origin chunking (bounded scratch / path memory) must not change a single bit."""
    flat, ref = util.load_golden("wind3d_small")
    ctx.set_model(flat)
    ctx.run(2)
    J1 = ctx.get_J()
    per_origin = flat["hnrays"] * flat["nfreqs"] * 16
    ctx.set_params(scratch_bytes=per_origin * 100, path_cache_bytes=1 << 22, cache_paths=0)
    st = ctx.run(2)
    assert st.n_chunks > 5
    assert np.array_equal(ctx.get_J(), J1)
    ctx.set_params(scratch_bytes=0, path_cache_bytes=0, cache_paths=1)


def test_origin_range_and_host_buffer_call(ctx):
    flat, ref = util.load_golden("wind3d_small")
    ctx.set_model(flat)
    N = flat["npoints"]
    J, (val, cnt), st = ctx.solve(flat["emissivity"], flat["opacity"], flat["inverse_width"], flat["population_tot"],
                                  approx=2, o_range=(N // 3, 2 * N // 3))
    assert util.relerr(J, ref["ref_J"][N // 3:2 * N // 3]) <= TOL
    assert util.relerr(val, ref["ref_lambda_val"][N // 3:2 * N // 3]) <= TOL


def test_errors_are_loud(ctx):
    from magritte_b200 import cabi
    c = cabi.Context(0)
    with pytest.raises(cabi.MgbError):
        c.npoints = 10
        c.run(2)
    with pytest.raises(cabi.MgbError):
        c.set_params(n_off_diag=5)  # more than kMaxOffDiag = 4 off-diagonals
    c.close()


@pytest.mark.parametrize("fn", ["expf", "logf", "erff", "expm1f", "powf"])
def test_device_libm_bit_exact(ctx, fn):
    """Device build of mgb_libm.h against the box's own libm on ~8 million inputs per function: a lattice over
    all float bit patterns plus samples dense in the ranges the path feeds (profile exponents, h nu / kT, ratios)."""
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library("m"))
    rng = np.random.default_rng(3)
    lattice = np.arange(0, 1 << 32, 1021, dtype=np.uint64).astype(np.uint32).view(np.float32)
    dense = {"expf": -rng.uniform(0, 110, 1 << 22), "logf": np.exp(rng.uniform(-80, 80, 1 << 22)),
             "erff": rng.uniform(-7, 7, 1 << 22), "expm1f": rng.uniform(0, 95, 1 << 22),
             "powf": np.exp(rng.uniform(-60, 60, 1 << 22))}[fn].astype(np.float32)
    x = np.concatenate([lattice, dense])
    y = None
    if fn == "powf":
        n = rng.integers(2, 40, x.size)
        y = (1.0 - (rng.integers(1, 40, x.size) % n) / n).astype(np.float32)
    got = ctx.libm(fn, x, y)
    f = getattr(libm, fn)
    f.restype = ctypes.c_float
    f.argtypes = [ctypes.c_float] * (2 if fn == "powf" else 1)
    idx = rng.choice(x.size, 200000, replace=False)
    want = np.array([f(float(x[i]), float(y[i])) if y is not None else f(float(x[i])) for i in idx], dtype=np.float32)
    g = got[idx]
    same = (g.view(np.uint32) == want.view(np.uint32)) | (np.isnan(g) & np.isnan(want))
    assert same.all(), f"{fn}: {np.count_nonzero(~same)} mismatches, first x={x[idx][~same][:3]}"


def test_ddiv_matches_ieee_division(ctx):
    """The kernels' scaled division (mgb_device.cuh: ddiv) must equal IEEE-754 round-to-nearest a/b bit for bit:
    16M random operand pairs over the whole exponent range (incl. the 1e-30..1e-60 magnitudes of SI emissivities),
    pairs with nearly equal mantissas, exact quotients, zeros, subnormals, inf and NaN. The self-test kernel also runs ddiv2
    (a[i] / b[i] and a[i^1] / b[i] with one reciprocal refinement) and poisons the output where its bits differ from ddiv's."""
    rng = np.random.default_rng(11)
    n = 1 << 24
    mant = lambda k: rng.integers(0, 1 << 52, k, dtype=np.uint64)
    expo = lambda k, lo, hi: rng.integers(lo, hi, k, dtype=np.uint64) << np.uint64(52)
    sign = lambda k: rng.integers(0, 2, k, dtype=np.uint64) << np.uint64(63)
    a = (sign(n) | expo(n, 1, 2047) | mant(n)).view(np.float64)
    b = (sign(n) | expo(n, 1, 2047) | mant(n)).view(np.float64)
    # typical magnitudes of the path
    a2 = np.exp(rng.uniform(-140, 10, n // 4)) * rng.choice([-1.0, 1.0], n // 4)
    b2 = np.exp(rng.uniform(-140, 10, n // 4))
    # hard-to-round: mantissas that differ in the last bits, quotients close to 1 and to powers of two
    m0 = mant(n // 4)
    a3 = (expo(n // 4, 900, 1100) | m0).view(np.float64)
    b3 = (expo(n // 4, 900, 1100) | (m0 + rng.integers(0, 4, n // 4, dtype=np.uint64))).view(np.float64)
    special = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308,
                        1e-310, 3.0, 1e300, 1e-300])
    a4, b4 = [x.ravel() for x in np.meshgrid(special, special)]
    # reciprocals (a == 1.0 is served by drcp, the unscaled Newton sequence): whole exponent range, typical magnitudes, specials
    nr = 1 << 22
    b5 = np.concatenate([(sign(nr) | expo(nr, 0, 2047) | mant(nr)).view(np.float64), np.exp(rng.uniform(-140, 140, nr)), special,
                         np.ldexp(1.0 + rng.integers(0, 4, 4096) * 2.0 ** -52, rng.integers(-1022, 1024, 4096))])
    # hard-to-round reciprocals: b = rn(1 / midpoint between two adjacent doubles) and its neighbours, mantissas of all ones / near 1
    nh = 1 << 21
    q = (rng.integers(1 << 52, 1 << 53, nh, dtype=np.uint64)).astype(np.longdouble)
    mid = (2 * q + 1) / np.longdouble(2.0) ** 54                      # exact in 64-bit-mantissa long double
    bh = (np.longdouble(1.0) / mid).astype(np.float64)
    bh = np.concatenate([bh, np.nextafter(bh, 0.0), np.nextafter(bh, 4.0)]) * np.ldexp(1.0, rng.integers(-60, 60, 3 * nh))
    ones_m = ((np.uint64(1) << np.uint64(52)) - rng.integers(1, 64, 4096, dtype=np.uint64)) | expo(4096, 900, 1100)
    near1 = rng.integers(0, 64, 4096, dtype=np.uint64) | expo(4096, 900, 1100)
    b6 = np.concatenate([bh, ones_m.view(np.float64), near1.view(np.float64)])
    A = np.concatenate([a, a2, a3, a4, np.ones(1000), rng.integers(1, 1 << 20, 1000).astype(np.float64), np.ones(b5.size), np.ones(b6.size)])
    B = np.concatenate([b, b2, b3, b4, rng.uniform(1, 2, 1000), rng.integers(1, 1 << 20, 1000).astype(np.float64), b5, b6])
    got = ctx.ddiv(A, B)
    with np.errstate(all="ignore"):
        want = A / B
    same = (got.view(np.uint64) == want.view(np.uint64)) | (np.isnan(got) & np.isnan(want))
    assert same.all(), f"{np.count_nonzero(~same)} mismatches, first: a={A[~same][:3]} b={B[~same][:3]} got={got[~same][:3]} want={want[~same][:3]}"


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_banded_lambda_against_reference(ctx, name):
    """n_off_diag > 0: J and the (p, line) Lambda lists of the compiled reference (same points, same list lengths, values within
    1e-10 after sorting each list by point), and the same against the oracle for the other ApproximationTypes."""
    from magritte_b200.flatten import approx_of
    from oracle import oracle as orc
    flat, ref = util.load_golden_offdiag(name)
    ctx.set_model(flat)
    approx = approx_of(flat["params"])
    ctx.run(approx)
    assert util.relerr(ctx.get_J(), ref["J"]) <= TOL
    val, nr, cnt = ctx.get_lambda_lists()
    cap = ref["lambda_nr"].shape[2]
    assert (nr[:, :, cap:] == 0xffffffff).all()
    util.lambda_lists_close(val[:, :, :cap], nr[:, :, :cap], cnt, ref, TOL)
    dval, dcnt = ctx.get_lambda()
    assert np.array_equal(dcnt, cnt) and np.array_equal(dval, val[:, :, 0])
    for other in (0, 1):
        o = orc.Oracle(flat, approx=other, store_u=False)
        o.solve()
        ctx.run(other)
        v2, n2, c2 = ctx.get_lambda_lists()
        k = o.L_nr.shape[2]
        want = dict(lambda_val=o.L_val, lambda_nr=o.L_nr, lambda_cnt=o.L_cnt)
        util.lambda_lists_close(v2[:, :, :k], n2[:, :, :k], c2, want, TOL)
    ctx.set_params(n_off_diag=0)
