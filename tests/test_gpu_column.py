"""Parity of mgb_set_column (k_column, through the C ABI) with Model::set_column (model.cpp:1143-1154 -> Solver::get_column,
solver.tpp:2968-2991): Model::column of the compiled reference (tests/golden/column_*.npz) bit for bit, and at a size the
reference was not run on the oracle plus the properties that do not depend on the size."""
import json
import os

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

NPOINTS = int(os.environ.get("MGB_COLUMN_POINTS", "100000"))  # configs[2] point count


@pytest.fixture(scope="module")
def ctx():
    from magritte_b200 import cabi
    c = cabi.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("name", util.GOLDEN_CASES)
def test_column_against_reference(ctx, name):
    flat, _ = util.load_golden(name)
    density, column = util.load_golden_column(name)
    ctx.set_model(flat)
    got = ctx.set_column(density)
    assert got.shape == column.shape
    assert np.array_equal(got, column), f"{name}: max rel err {util.relerr(got, column)}"
    lo, hi = 1, min(int(flat["npoints"]), 23)
    assert np.array_equal(ctx.set_column(density, o_range=(lo, hi)), column[:, lo:hi])


def test_column_errors(ctx):
    from magritte_b200 import cabi
    flat, _ = util.load_golden("c1_all_constant")
    density, _ = util.load_golden_column("c1_all_constant")
    ctx.set_model(flat)
    for rng in [(5, 5), (7, 3), (0, int(flat["npoints"]) + 1)]:
        with pytest.raises(cabi.MgbError):
            ctx.set_column(density, o_range=rng)
    fresh = cabi.Context(0)
    try:
        with pytest.raises(cabi.MgbError):
            fresh.npoints, fresh.nrays = 4, 2
            fresh.set_column(np.ones(4))
    finally:
        fresh.close()


def test_column_large_properties():
    """configs[2]-shaped point cloud (3D wind, HEALPix nside 4: 192 rays): oracle rows of two blocks of origins bit for bit;
    scaling the density by a power of two scales every column exactly; unit density gives the geometric path length, which is the
    same seen from both ends of a ray for boundary-to-boundary chords through the origin (Z sums differ only by rounding)."""
    from magritte_b200 import cabi, modelbuild as mb, synth
    from oracle import oracle as orc
    spec = mb.spec_wind_3d(npoints=NPOINTS, nside_rays=4, nquads=3, nlines_keep=range(2), seed=4321)
    flat = synth.flat_from_spec(spec)
    N, R = int(flat["npoints"]), int(flat["nrays"])
    rng = np.random.default_rng(99)
    density = 10.0 ** rng.uniform(6.0, 14.0, size=N)
    c = cabi.Context(0)
    try:
        c.set_model(flat)
        col, ms = c.set_column(density, with_time=True)
        col4 = c.set_column(4.0 * density)
        length = c.set_column(np.ones(N))
    finally:
        c.close()
    assert col.shape == (R, N) and np.all(np.isfinite(col)) and np.all(col >= 0.0)
    assert np.array_equal(col4, 4.0 * col)
    o = orc.Oracle(flat, store_u=False, threads=os.cpu_count() or 1)
    for start in (N // 5, N - 2048):
        want = o.set_column(density, o_range=(start, start + 2048))
        assert np.array_equal(col[:, start:start + 2048], want)
    # column density is bounded by max density x path length
    assert np.all(col <= density.max() * length * (1 + 1e-12))
    print(json.dumps({"column": {"npoints": N, "nrays": R, "kernel_ms": ms, "half_rays_per_s": N * R / (ms * 1e-3)}}))


@pytest.mark.parametrize("name", ["c2_vanzadelhoff", "wind3d_small"])
def test_column_through_host_model(name):
    """magritte_b200.core.Model: density -> set_column() -> column, the reference's attribute and method names (pybindings.cpp:138-139, 268)."""
    from test_host_model import host_model
    m, flat, ref, lp = host_model(name)
    density, column = util.load_golden_column(name)
    with pytest.raises(RuntimeError):
        m.set_column()  # no density yet
    m.density = density
    assert m.set_column() == 0
    assert np.array_equal(m.column, column)
