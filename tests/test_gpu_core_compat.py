"""The drop-in boundary: ONE script written against `magritte.core` -- this repository's own driver `modelbuild.build()` +
`prepare()`, the code that drives the compiled reference everywhere else in the suite -- is run unchanged against (a) the compiled
reference (oracle/_ref/magritte/core.so) and (b) `magritte_b200.core_compat`, the reference's Python surface over the B200 host layer
(attribute tree, `.set(ndarray)` containers, `Model.write(io)` / `read(io)` over an in-memory Io, the compute_* overloads). The two
must produce the same J, Lambda lists, level populations and images."""
import numpy as np
import pytest

import util
from test_host_model import assert_populations_close

pytestmark = pytest.mark.gpu
CASES = ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small", "wind3d_spiral_2spec"]


def both(name):
    from magritte_b200 import core_compat, modelbuild as mb
    from oracle import refcore
    if not refcore.available():
        pytest.skip("oracle/_ref not built")
    core_ref, _ = refcore.load()
    core_ref.pcmt_set_n_threads_avail(8)  # as in tests/golden/make_golden*.py
    spec = util.small_specs()[name]
    out = []
    for core in (core_ref, core_compat):
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", f"compat_{name}"), f"compat_{name}", spec)
            mb.prepare(m)
        out.append(m)
    return spec, out[0], out[1]


def lambda_lists(m):
    out = []
    for ls in m.lines.lineProducingSpecies:
        out.append(([[list(map(float, x)) for x in p] for p in ls.Lambda.Ls], [[list(map(int, x)) for x in p] for p in ls.Lambda.nr]))
    return out


@pytest.mark.parametrize("name", CASES)
def test_same_script_same_radiation_field(name):
    spec, ref, new = both(name)
    # what read() derives and prepare() computes, bit for bit
    assert np.array_equal(np.array(ref.geometry.rays.antipod), np.array(new.geometry.rays.antipod))
    assert np.array_equal(np.array(ref.geometry.boundary.point2boundary), np.array(new.geometry.boundary.point2boundary))
    assert np.array_equal(np.array(ref.lines.line), np.array(new.lines.line))
    assert ref.parameters.nlines() == new.parameters.nlines() and ref.parameters.hnrays() == new.parameters.hnrays()
    assert np.array_equal(np.array(ref.radiation.frequencies.nu), np.array(new.radiation.frequencies.nu))
    assert np.array_equal(np.array(ref.lines.inverse_width), np.array(new.lines.inverse_width))
    assert np.array_equal(np.array(ref.lines.emissivity), np.array(new.lines.emissivity))
    assert np.array_equal(np.array(ref.lines.opacity), np.array(new.lines.opacity))
    for a, b in zip(ref.lines.lineProducingSpecies, new.lines.lineProducingSpecies):
        assert np.array_equal(np.array(a.population, dtype=np.float64), np.array(b.population, dtype=np.float64))
    with util.quiet():
        assert ref.compute_radiation_field_feautrier_order_2() == 0
    assert new.compute_radiation_field_feautrier_order_2() == 0
    assert util.relerr(np.array(new.radiation.J), np.array(ref.radiation.J)) <= 1e-10
    for (Ls_r, nr_r), (Ls_n, nr_n) in zip(lambda_lists(ref), lambda_lists(new)):
        assert nr_r == nr_n
        flat_r = np.array([v for p in Ls_r for k in p for v in k])
        flat_n = np.array([v for p in Ls_n for k in p for v in k])
        assert util.relerr(flat_n, flat_r) <= 1e-10


@pytest.mark.parametrize("name", ["c1_all_constant", "c2_vanzadelhoff", "wind3d_small"])
def test_same_script_same_level_populations(name):
    spec, ref, new = both(name)
    K = 3
    for m in (ref, new):
        m.parameters.use_adaptive_Ng_acceleration = False
    with util.quiet():
        assert ref.compute_level_populations(False, K) == K
    assert new.compute_level_populations(False, K) == K
    for a, b in zip(ref.lines.lineProducingSpecies, new.lines.lineProducingSpecies):
        nlev = a.linedata.nlev
        got = np.array(b.population, dtype=np.float64).reshape(-1, nlev)
        assert_populations_close(got, np.array(a.population, dtype=np.float64), np.array(a.population_tot), name)
        assert util.relerr(np.array(b.Jeff), np.array(a.Jeff)) <= 1e-10
    assert util.relerr(np.array(new.radiation.J), np.array(ref.radiation.J)) <= 1e-10
    assert np.allclose(new.error_max, ref.error_max, rtol=1e-8, atol=1e-14)


@pytest.mark.parametrize("name", ["wind3d_small", "wind3d_spiral_2spec"])
def test_same_script_same_image(name):
    """The imaging calls of the script -- compute_spectral_discretisation(nu_min, nu_max, n), compute_image_new(rx, ry, rz, Nx, Ny) and the
    (ray_nr, Nx, Ny) overload -- through the facade, against the images the compiled reference produced for exactly these calls
    (tests/golden/image_<name>.npz, made by make_golden_image.py with the same driver script): ImX, ImY and I bit for bit.
    The reference is not run live here: its threaded imager segfaulted now and then inside the long-lived pytest process on the GPU box
    (never in a process of its own, tools/image_crash_probe.py), which took the whole suite down."""
    import os
    from magritte_b200 import core_compat, modelbuild as mb
    sys_path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    import importlib.util
    mod = importlib.util.spec_from_file_location("make_golden_image", os.path.join(sys_path, "make_golden_image.py"))
    cases = importlib.util.module_from_spec(mod)
    mod.loader.exec_module(cases)
    dim, li, dv, nf, dirs, (nx, ny) = cases.CASES[name]
    g = np.load(os.path.join(sys_path, f"image_{name}.npz"))
    spec = util.small_specs()[name]
    with util.quiet():
        m = mb.build(core_compat, core_compat.IoPython("mem", f"compat_img_{name}"), f"compat_img_{name}", spec)
        mb.prepare(m)
    if name == "wind3d_spiral_2spec":
        m.parameters.sum_opacity_emissivity_over_all_lines = False
    nu0 = float(np.array(m.lines.line)[li])
    assert m.compute_spectral_discretisation(nu0 * (1 - dv / 2.99792458e8), nu0 * (1 + dv / 2.99792458e8), nf) == 0
    for i, d in enumerate(dirs):
        assert m.compute_image_new(float(d[0]), float(d[1]), float(d[2]), nx, ny) == 0   # (rx, ry, rz, Nxpix, Nypix)
        im = m.images[-1]
        assert np.array_equal(np.array(im.ImX), g[f"ImX_{i}"]) and np.array_equal(np.array(im.ImY), g[f"ImY_{i}"])
        assert np.array_equal(np.array(im.I), g[f"I_{i}"])
    # (ray_nr, Nxpix, Nypix): the direction of ray ray_nr (model.cpp:985-997)
    d1 = np.array(m.geometry.rays.direction)[1]
    assert m.compute_image_new(1, 16, 12) == 0
    assert m.compute_image_new(float(d1[0]), float(d1[1]), float(d1[2]), 16, 12) == 0
    a, b = m.images[-2], m.images[-1]
    assert np.array_equal(np.array(a.ImX), np.array(b.ImX)) and np.array_equal(np.array(a.I), np.array(b.I))


def test_container_and_parameter_behaviour():
    """No GPU needed for this part, but it lives with the boundary tests: `.set()` checks dtype and rank like the pybind containers,
    essentials are set-once, a missing model fails like Model::read."""
    from magritte_b200 import core_compat as core
    m = core.Model()
    with pytest.raises(TypeError):
        m.geometry.points.neighbors.set(np.zeros(4, dtype=np.int64))
    with pytest.raises(RuntimeError):
        m.geometry.points.position.set(np.zeros(3))
    m.parameters.set_npoints(5)
    m.parameters.set_npoints(5)
    with pytest.raises(RuntimeError, match="SetOnce"):
        m.parameters.set_npoints(6)
    with pytest.raises(RuntimeError, match="SetOnce"):
        m.parameters.nrays()
    m2 = core.Model()
    m2.parameters.set_model_name("no_such_model")
    with pytest.raises(OSError):
        m2.read(core.IoPython("mem", "no_such_model"))
    assert core.BoundaryCondition.CMB == 2 and len(core.vCollisionPartner([core.CollisionPartner()])) == 1
