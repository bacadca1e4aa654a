"""mgb_set_column_noalloc (magritte_b200/csrc/mgb_noalloc.cu): Model::set_column staged through the library's file-scope static device
array instead of per-call device buffers -- an entry point that allocates nothing on the device, for callers that cannot call cudaMalloc.
It has no counterpart in the reference; its yardstick is the existing entry point mgb_set_column, whose results it must reproduce
bit for bit (the tolerance of tests/test_gpu_column.py: np.array_equal), on calls in a row that share the array.

Collected last on purpose: it passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

MAX_DOUBLES = 4 << 20  # kNoAllocMaxDoubles in mgb_noalloc.cu, stated in include/magritte_b200.h


def test_noalloc_column_matches_set_column_on_calls_in_a_row():
    from magritte_b200 import cabi
    flat, _ = util.load_golden("wind3d_small")
    density, column = util.load_golden_column("wind3d_small")
    N = int(flat["npoints"])
    rng = np.random.default_rng(5)
    other = 10.0 ** rng.uniform(3.0, 9.0, size=N)
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        want1 = ctx.set_column(density)
        want2 = ctx.set_column(other)
        got1 = ctx.set_column_noalloc(density)           # first call: fetches the array's device address
        got2 = ctx.set_column_noalloc(other)             # second call in a row: finds the first call's contents in the array
        got3 = ctx.set_column_noalloc(density, (N // 3, N // 2))  # a smaller block over the stale larger one
        got4 = ctx.set_column_noalloc(density)
    finally:
        ctx.close()
    assert np.array_equal(want1, column)                 # the yardstick itself is the reference's Model::column
    assert got1.shape == want1.shape and np.array_equal(got1, want1)
    assert np.array_equal(got2, want2)
    assert np.array_equal(got3, want1[:, N // 3:N // 2])
    assert np.array_equal(got4, want1)


def test_noalloc_column_is_shared_by_contexts_and_checks_its_arguments():
    """A second context on the same device uses the same array (sequentially: fine); bad ranges and a missing geometry are errors."""
    from magritte_b200 import cabi
    flat, _ = util.load_golden("c2_vanzadelhoff")
    density, column = util.load_golden_column("c2_vanzadelhoff")
    a, b = cabi.Context(0), cabi.Context(0)
    try:
        a.set_model(flat)
        b.set_model(flat)
        assert np.array_equal(a.set_column_noalloc(density), column)
        assert np.array_equal(b.set_column_noalloc(2.0 * density), 2.0 * column)
        assert np.array_equal(a.set_column_noalloc(density), column)
        for rng in [(5, 5), (7, 3), (0, int(flat["npoints"]) + 1)]:
            with pytest.raises(cabi.MgbError):
                a.set_column_noalloc(density, o_range=rng)
    finally:
        a.close()
        b.close()
    fresh = cabi.Context(0)
    try:
        fresh.npoints, fresh.nrays = 4, 2
        with pytest.raises(cabi.MgbError):
            fresh.set_column_noalloc(np.ones(4))
    finally:
        fresh.close()


def test_noalloc_column_rejects_sizes_beyond_the_static_array():
    """40 000 points x 192 rays = 7.7 M doubles > 4 Mi: the full range is rejected with the sizes in the message; a range that fits
    (npoints + 192 * origins <= 4 Mi) is served and equals the allocating entry point's rows."""
    from magritte_b200 import cabi, modelbuild as mb, synth
    flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=40000, nside_rays=4, nquads=3, nlines_keep=range(2), seed=77))
    N, R = int(flat["npoints"]), int(flat["nrays"])
    assert N + N * R > MAX_DOUBLES
    fit = (MAX_DOUBLES - (N + 31) // 32 * 32) // R
    density = 10.0 ** np.random.default_rng(3).uniform(6.0, 12.0, size=N)
    ctx = cabi.Context(0)
    try:
        ctx.set_model(flat)
        with pytest.raises(cabi.MgbError, match="exceed the static scratch"):
            ctx.set_column_noalloc(density)
        with pytest.raises(cabi.MgbError, match="exceed the static scratch"):
            ctx.set_column_noalloc(density, (0, fit + 1))
        got = ctx.set_column_noalloc(density, (0, fit))
        want = ctx.set_column(density, (0, fit))
        tail = ctx.set_column_noalloc(density, (N - 100, N))
        want_tail = ctx.set_column(density, (N - 100, N))
    finally:
        ctx.close()
    assert np.array_equal(got, want) and np.array_equal(tail, want_tail)
