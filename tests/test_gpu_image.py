"""Parity of the CUDA imager (mgb_image_feautrier_order_2, through the C ABI) with Solver::image_feautrier_order_2_new_imager:
Image.ImX / ImY / I of the compiled reference (tests/golden/image_*.npz) and of the oracle. The imager keeps the reference's
forward elimination order, so the bar here is bit-exactness, not only the 1e-10 of BASELINE.json's north_star."""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from magritte_b200 import cabi
    c = cabi.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_image_against_reference(ctx, name):
    from magritte_b200.flatten import approx_of
    flat, img = util.load_golden_image(name)
    ctx.set_model(flat, frequencies=False)
    approx = approx_of(flat["params"])
    for d, (X, Y, I) in zip(img["directions"], img["images"]):
        gX, gY, gI, st = ctx.image(approx, img["dimension"], d, img["nx"], img["ny"], flat["nu_image"])
        assert np.array_equal(gX, X) and np.array_equal(gY, Y)
        assert util.relerr(gI, I) <= 1e-10, f"{name} {d}"
        assert np.array_equal(gI, I), f"{name} {d}: not bit-identical, max rel err {util.relerr(gI, I)}"
        assert st.n_solves == I.size


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_optical_depth_image_against_reference(ctx, name):
    """mgb_image_optical_depth against Model::compute_image_optical_depth_new of the compiled reference (same summation order)."""
    from magritte_b200.flatten import approx_of
    flat, img = util.load_golden_image(name)
    ctx.set_model(flat, frequencies=False)
    approx = approx_of(flat["params"])
    for d, (X, Y, T) in zip(img["directions"], util.load_golden_image_tau(name)):
        gX, gY, gT, st = ctx.image(approx, img["dimension"], d, img["nx"], img["ny"], flat["nu_image"], optical_depth=True)
        assert np.array_equal(gX, X) and np.array_equal(gY, Y)
        assert util.relerr(gT, T) <= 1e-10, f"{name} {d}"
        assert np.array_equal(gT, T), f"{name} {d}: not bit-identical, max rel err {util.relerr(gT, T)}"


def test_image_options_and_errors(ctx):
    """Per-point frequency rows, pixel sub-ranges, small chunks (bounded storage) and ApproximationType None give the same
    image; the reference's exceptions map to error codes."""
    from magritte_b200 import cabi
    from oracle import oracle as orc
    flat, img = util.load_golden_image("wind3d_small")
    ctx.set_model(flat, frequencies=False)
    d, (X, Y, I) = img["directions"][1], img["images"][1]
    nx, ny = img["nx"], img["ny"]
    _, _, I2, _ = ctx.image(2, 3, d, nx, ny, flat["nu"])  # [npoints][nfreqs]
    assert np.array_equal(I2, I)
    _, _, I3, _ = ctx.image(2, 3, d, nx, ny, flat["nu_image"], pix_range=(37, 301))
    assert np.array_equal(I3, I[37:301])
    ctx.set_params(path_cache_bytes=1, scratch_bytes=8 * 31 * 50)
    _, _, I4, st = ctx.image(2, 3, d, nx, ny, flat["nu_image"])
    assert st.n_chunks > 1 and np.array_equal(I4, I)
    ctx.set_params(path_cache_bytes=0, scratch_bytes=0)
    want = orc.Oracle(flat, approx=0, store_u=False).image(d, nx, ny, 3)["I"]
    _, _, I5, _ = ctx.image(0, 3, d, nx, ny, flat["nu_image"])
    assert np.array_equal(I5, want)
    for args in [(1, 3, d), (2, 3, (0.0, 0.0, 0.0)), (2, 1, (1.0, 0.0, 0.0)), (2, 2, d)]:
        with pytest.raises(cabi.MgbError):
            ctx.image(args[0], args[1], args[2], nx, ny, flat["nu_image"])


@pytest.mark.parametrize("name", ["velgrad_1d", "wind3d_spiral_2spec"])
def test_image_kernels_bit_identical(ctx, name, monkeypatch):
    """The record path (single-pass trace + k_image_records + k_image_solve_rec, the default), its fallback when a walk does not
    fit the per-pixel capacity, and the generic kernel (MGB_IMAGE_KERNEL=1) give the same bits, intensities and optical depths."""
    from magritte_b200.flatten import approx_of
    flat, img = util.load_golden_image(name)
    ctx.set_model(flat, frequencies=False)
    approx = approx_of(flat["params"])
    d = img["directions"][-1]
    for tau in (False, True):
        monkeypatch.delenv("MGB_IMAGE_KERNEL", raising=False)
        monkeypatch.delenv("MGB_IMAGE_CAP", raising=False)
        _, _, I0, st0 = ctx.image(approx, img["dimension"], d, img["nx"], img["ny"], flat["nu_image"], optical_depth=tau)
        monkeypatch.setenv("MGB_IMAGE_CAP", "3")
        _, _, I1, st1 = ctx.image(approx, img["dimension"], d, img["nx"], img["ny"], flat["nu_image"], optical_depth=tau)
        monkeypatch.delenv("MGB_IMAGE_CAP")
        monkeypatch.setenv("MGB_IMAGE_KERNEL", "1")
        _, _, I2, st2 = ctx.image(approx, img["dimension"], d, img["nx"], img["ny"], flat["nu_image"], optical_depth=tau)
        assert np.array_equal(I0, I1) and np.array_equal(I0, I2)
        assert st0.n_entries == st1.n_entries == st2.n_entries and st0.max_ray_length == st2.max_ray_length
        if st0.max_ray_length > 3:
            assert st1.n_launches == st0.n_launches + 1  # the second trace pass
