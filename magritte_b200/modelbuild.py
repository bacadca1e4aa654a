"""Model construction helpers and the synthetic workloads of BASELINE.json.

Host-side pre-processing in the spirit of the reference's `magritte.setup`
(reference: magritte/setup.py), which cannot be imported in this image (needs h5py,
healpy, astroquery). Everything here only talks to the `magritte.core` API surface
(`Model`, `Parameters` setters, `.set(ndarray)` containers), so the same builder drives
either the compiled reference (oracle/_ref, used as the parity oracle) or this
repository's own `magritte_b200.core`.

Workloads:
  * all_constant_1d          -- BASELINE config 0 (tests/benchmarks/analytic/all_constant_single_ray.py:16-65)
  * vanzadelhoff_1a_1d       -- BASELINE config 1 (tests/benchmarks/numeric/vanZadelhoff_1_1D.py:19-83)
  * wind_3d                  -- BASELINE configs 2/3: synthetic 3D wind point cloud (seeded)
"""
import math
import os
import numpy as np

from . import lamda
from .lamda import CC, HH, KB, AMU, T_CMB

_HERE = os.path.dirname(os.path.abspath(__file__))
DATA_DIR = os.path.join(os.path.dirname(_HERE), "tests", "golden")


# --------------------------------------------------------------------------------------
# HEALPix (ring ordering) pixel centres -- replaces healpy.pixelfunc.pix2vec
# --------------------------------------------------------------------------------------
def healpix_pix2vec_ring(nside):
    """Unit vectors of all 12*nside^2 HEALPix pixel centres in RING ordering."""
    npix = 12 * nside * nside
    ncap = 2 * nside * (nside - 1)
    vec = np.zeros((npix, 3))
    for p in range(npix):
        if p < ncap:
            i = (1 + math.isqrt(1 + 2 * p)) >> 1
            j = (p + 1) - 2 * i * (i - 1)
            z = 1.0 - (i * i) / (3.0 * nside * nside)
            phi = (j - 0.5) * math.pi / (2.0 * i)
        elif p < npix - ncap:
            ip = p - ncap
            i = ip // (4 * nside) + nside
            j = ip % (4 * nside) + 1
            fodd = 0.5 * (1 + ((i + nside) & 1))
            z = (2 * nside - i) * 2.0 / (3.0 * nside)
            phi = (j - fodd) * math.pi / (2.0 * nside)
        else:
            ip = npix - p
            i = (1 + math.isqrt(2 * ip - 1)) >> 1
            j = 4 * i + 1 - (ip - 2 * i * (i - 1))
            z = -1.0 + (i * i) / (3.0 * nside * nside)
            phi = (j - 0.5) * math.pi / (2.0 * i)
        s = math.sqrt(max(0.0, (1.0 - z) * (1.0 + z)))
        vec[p] = (s * math.cos(phi), s * math.sin(phi), z)
    return vec


def antipodal_ray_set(nside):
    """HEALPix directions re-ordered so that ray r and ray r+hnrays are exact antipodes.

    The reference solver loops over rr < hnrays and pairs it with antipod[rr]
    (reference: src/solver/solver.tpp:476-479), so the first half must hold one member of
    every antipodal pair. The second half is the exact negation (bit-for-bit), which makes
    the antipode search of reference src/model/geometry/rays/rays.cpp:52-63 unambiguous.
    """
    v = healpix_pix2vec_ring(nside)
    h = v.shape[0] // 2
    first = v[:h]
    return np.concatenate([first, -first], axis=0)


# --------------------------------------------------------------------------------------
# Pieces of magritte.setup restated
# --------------------------------------------------------------------------------------
def neighbors_1d(npoints):
    """reference: magritte/setup.py:59-70"""
    nbs = [1]
    n_nbs = [1]
    for p in range(1, npoints - 1):
        nbs.extend([p - 1, p + 1])
        n_nbs.append(2)
    nbs.append(npoints - 2)
    n_nbs.append(1)
    return np.array(nbs, dtype=np.uint32), np.array(n_nbs, dtype=np.uint32)


def neighbors_delaunay(position):
    """reference: magritte/setup.py:80-88"""
    from scipy.spatial import Delaunay
    d = Delaunay(np.asarray(position))
    indptr, nbs = d.vertex_neighbor_vertices
    n_nbs = np.diff(indptr)
    return np.asarray(nbs, dtype=np.uint32), np.asarray(n_nbs, dtype=np.uint32)


def rays_spherical_symmetry(nrays):
    """Uniform ray set for 1D spherical symmetry. reference: magritte/setup.py:226-339 (uniform=True)."""
    nextra = nrays // 2 - 1
    Rx, Ry, Rz = [1.0], [0.0], [0.0]
    angle_max = np.arctan(Ry[-1] / Rx[-1])
    angle_extra = (0.5 * np.pi - angle_max) / nextra
    for k in range(1, nextra):
        Rx.append(np.cos(angle_max + k * angle_extra))
        Ry.append(np.sin(angle_max + k * angle_extra))
        Rz.append(0.0)
    Rx.append(0.0)
    Ry.append(1.0)
    Rz.append(0.0)
    n2 = len(Rx)
    Wt = []
    for n in range(n2):
        if n == 0:
            ux, uy = 0.5 * (Rx[n] + Rx[n + 1]), 0.5 * (Ry[n] + Ry[n + 1])
            lx, ly = Rx[0], Ry[0]
        elif n == n2 - 1:
            ux, uy = Rx[-1], Ry[-1]
            lx, ly = 0.5 * (Rx[n] + Rx[n - 1]), 0.5 * (Ry[n] + Ry[n - 1])
        else:
            ux, uy = 0.5 * (Rx[n] + Rx[n + 1]), 0.5 * (Ry[n] + Ry[n + 1])
            lx, ly = 0.5 * (Rx[n] + Rx[n - 1]), 0.5 * (Ry[n] + Ry[n - 1])
        Wt.append(lx / np.sqrt(lx**2 + ly**2) - ux / np.sqrt(ux**2 + uy**2))
    for n in range(n2):
        Rx.append(-Rx[n])
        Ry.append(-Ry[n])
        Rz.append(-Rz[n])
        Wt.append(Wt[n])
    direction = np.array((Rx, Ry, Rz)).transpose()
    weight = np.array(Wt)
    weight /= np.sum(weight)
    return direction, weight


def gauss_hermite(nquads):
    """reference: magritte/setup.py:480-503"""
    roots, weights = np.polynomial.hermite.hermgauss(nquads)
    weights = weights / np.sum(weights)
    return roots, weights


def set_linedata(core, model, lspec, ld, symbols):
    """Fill model.lines.lineProducingSpecies[lspec].linedata from a parsed LAMDA dict.
    reference: magritte/setup.py:660-776"""
    L = model.lines.lineProducingSpecies[lspec].linedata
    L.sym = ld["sym"]
    L.num = symbols.index(ld["sym"])
    L.inverse_mass = float(ld["inverse_mass"])
    L.nlev = int(ld["nlev"])
    L.energy = [float(x) for x in ld["energy"]]
    L.weight = [float(x) for x in ld["weight"]]
    ncolpar = len(ld["colpar"])
    L.ncolpar = ncolpar
    L.colpar = core.vCollisionPartner([core.CollisionPartner() for _ in range(ncolpar)])
    for c, cp in enumerate(ld["colpar"]):
        P = L.colpar[c]
        P.num_col_partner = symbols.index(cp["name"]) if cp["name"] in symbols else 0
        P.orth_or_para_H2 = cp["orth_or_para_H2"]
        P.ncol = int(cp["ncol"])
        P.ntmp = int(cp["ntmp"])
        P.icol = [int(x) for x in cp["icol"]]
        P.jcol = [int(x) for x in cp["jcol"]]
        P.tmp = [float(x) for x in cp["tmp"]]
        P.Cd = [[float(x) for x in row] for row in cp["Cd"]]
        P.Ce = [[float(x) for x in row] for row in cp["Ce"]]
    L.nrad = int(ld["nrad"])
    L.irad = [int(x) for x in ld["irad"]]
    L.jrad = [int(x) for x in ld["jrad"]]
    L.A = [float(x) for x in ld["A"]]
    L.Bs = [float(x) for x in ld["Bs"]]
    L.Ba = [float(x) for x in ld["Ba"]]
    L.frequency = [float(x) for x in ld["frequency"]]


def restrict_lines(ld, keep):
    """Keep only radiative transitions `keep` (reference: setup.py:741-750 'considered transitions')."""
    out = dict(ld)
    keep = list(keep)
    out["nrad"] = len(keep)
    for k in ("irad", "jrad", "A", "Bs", "Ba", "frequency"):
        out[k] = np.asarray(ld[k])[keep]
    return out


def scaled_species(ld, sym, mass, energy_scale, A_scale):
    """A deterministic second species derived from `ld` (SURVEY 8d, C4 'SiO-like')."""
    out = dict(ld)
    out["sym"] = sym
    out["mass"] = mass
    out["inverse_mass"] = float(1.0 / mass)
    out["energy"] = ld["energy"] * energy_scale
    irad, jrad = ld["irad"], ld["jrad"]
    out["frequency"] = (out["energy"][irad] - out["energy"][jrad]) / HH
    out["A"] = ld["A"] * A_scale * energy_scale**3
    out["Bs"] = out["A"] * CC**2 / (2.0 * HH * out["frequency"]**3)
    out["Ba"] = ld["weight"][irad] / ld["weight"][jrad] * out["Bs"]
    cps = []
    for cp in ld["colpar"]:
        c2 = dict(cp)
        Ce = np.zeros_like(cp["Cd"])
        for t_ in range(cp["ntmp"]):
            Ce[t_] = cp["Cd"][t_] * ld["weight"][cp["icol"]] / ld["weight"][cp["jcol"]] * np.exp(
                -(out["energy"][cp["icol"]] - out["energy"][cp["jcol"]]) / (KB * cp["tmp"][t_]))
        c2["Ce"] = Ce
        cps.append(c2)
    out["colpar"] = cps
    return out


def load_species(name):
    return lamda.load_linedata(os.path.join(DATA_DIR, f"lamda_{name}.npz"))


# --------------------------------------------------------------------------------------
# Generic assembly through the magritte.core API
# --------------------------------------------------------------------------------------
def assemble(core, io, name, *, dimension, spherical, position, velocity, nbs, n_nbs, boundary2point,
             direction, weight, symbols, abundance, temperature, vturb2, linedata, nquads,
             boundary_condition="CMB", boundary_temperature=None, use_adaptive_directions=False):
    """Build a model through the pybind setters, write it to `io`, read it back (required:
    reference Lines::read builds the derived line tables, src/model/lines/lines.cpp:11-69) and
    return the freshly read model."""
    npoints = len(position)
    m = core.Model()
    P = m.parameters
    P.set_spherical_symmetry(bool(spherical))
    P.set_model_name(name)
    P.set_dimension(int(dimension))
    P.set_npoints(int(npoints))
    P.set_nrays(int(len(direction)) // (npoints if use_adaptive_directions else 1))  # adaptive: direction is [npoints * nrays][3]
    P.set_nspecs(len(symbols))
    P.set_nlspecs(len(linedata))
    P.set_nquads(int(nquads))
    P.set_nboundary(int(len(boundary2point)))
    m.geometry.points.position.set(np.ascontiguousarray(position, dtype=np.float64))
    m.geometry.points.velocity.set(np.ascontiguousarray(velocity, dtype=np.float64))
    m.geometry.points.neighbors.set(np.ascontiguousarray(nbs, dtype=np.uint32))
    m.geometry.points.n_neighbors.set(np.ascontiguousarray(n_nbs, dtype=np.uint32))
    m.chemistry.species.abundance = np.asarray(abundance, dtype=np.float64).tolist()
    m.chemistry.species.symbol = list(symbols)
    m.thermodynamics.temperature.gas.set(np.ascontiguousarray(temperature, dtype=np.float64))
    m.thermodynamics.turbulence.vturb2.set(np.ascontiguousarray(vturb2, dtype=np.float64))
    m.geometry.boundary.boundary2point.set(np.ascontiguousarray(boundary2point, dtype=np.uint32))
    bc = getattr(core.BoundaryCondition, boundary_condition)
    for b in range(len(boundary2point)):
        m.geometry.boundary.set_boundary_condition(b, bc)
    if boundary_temperature is None:
        boundary_temperature = np.full(len(boundary2point), T_CMB)
    m.geometry.boundary.boundary_temperature.set(np.ascontiguousarray(boundary_temperature, dtype=np.float64))
    m.geometry.rays.direction.set(np.ascontiguousarray(direction, dtype=np.float64))
    m.geometry.rays.weight.set(np.ascontiguousarray(weight, dtype=np.float64))
    m.geometry.rays.use_adaptive_directions = bool(use_adaptive_directions)
    m.lines.resize_LineProducingSpecies(len(linedata))
    roots, weights = gauss_hermite(nquads)
    for l, ld in enumerate(linedata):
        set_linedata(core, m, l, ld, list(symbols))
        m.lines.lineProducingSpecies[l].quadrature.roots.set(roots)
        m.lines.lineProducingSpecies[l].quadrature.weights.set(weights)
    m.write(io)
    m2 = core.Model()
    m2.parameters.set_model_name(name)
    m2.read(io)
    return m2


def prepare(model):
    """The standard post-read sequence of every reference test (e.g. all_constant_single_ray.py:84-88)."""
    model.compute_spectral_discretisation()
    model.compute_inverse_line_widths()
    model.compute_LTE_level_populations()
    return model


# --------------------------------------------------------------------------------------
# Workloads
# --------------------------------------------------------------------------------------
def spec_all_constant_1d(npoints=50, nquads=1):
    """BASELINE config 0. reference: tests/benchmarks/analytic/all_constant_single_ray.py:16-65"""
    nH2, nTT, temp, turb, dx = 1.0E+12, 1.0E+03, 4.5, 0.0, 1.0E+12
    position = np.array([[i * dx, 0, 0] for i in range(npoints)], dtype=np.float64)
    velocity = np.zeros((npoints, 3))
    nbs, n_nbs = neighbors_1d(npoints)
    return dict(dimension=1, spherical=False, position=position, velocity=velocity, nbs=nbs, n_nbs=n_nbs,
                boundary2point=np.array([0, npoints - 1], dtype=np.uint32),
                direction=np.array([[-1.0, 0, 0], [1.0, 0, 0]]), weight=np.array([0.5, 0.5]),
                symbols=["test", "H2", "e-"], abundance=np.array([[nTT, nH2, 0.0]] * npoints),
                temperature=temp * np.ones(npoints), vturb2=(turb / CC)**2 * np.ones(npoints),
                linedata=[load_species("test")], nquads=nquads)


def spec_vanzadelhoff_1a_1d(npoints=250, nrays=200, nquads=11, X_mol=1.0E-8):
    """BASELINE config 1. reference: tests/benchmarks/numeric/vanZadelhoff_1_1D.py:19-83"""
    r_in, r_out, nH2_in, temp, turb = 1.0E13, 7.8E16, 2.0E13, 20.0, 150.0
    rs = np.logspace(np.log10(r_in), np.log10(r_out), npoints, endpoint=True)
    nH2 = nH2_in * np.power(r_in / rs, 2.0)
    nTT = X_mol * nH2
    position = np.array([[r, 0, 0] for r in rs])
    nbs, n_nbs = neighbors_1d(npoints)
    direction, weight = rays_spherical_symmetry(nrays)
    return dict(dimension=1, spherical=True, position=position, velocity=np.zeros((npoints, 3)),
                nbs=nbs, n_nbs=n_nbs, boundary2point=np.array([0, npoints - 1], dtype=np.uint32),
                direction=direction, weight=weight, symbols=["test", "H2", "e-"],
                abundance=np.stack([nTT, nH2, np.zeros(npoints)], axis=1),
                temperature=temp * np.ones(npoints), vturb2=(turb / CC)**2 * np.ones(npoints),
                linedata=[load_species("test")], nquads=nquads)


def spec_velocity_gradient_1d(npoints=60, nrays=40, nquads=9, v_out=3.0e4):
    """Spherically symmetric model WITH Doppler shifts (pattern of
    tests/benchmarks/analytic/constant_velocity_gradient_1D.py:19-80); exercises the
    spherical get_shift branch and the large-shift optical depth."""
    r_in, r_out, nH2_in, temp, turb = 1.0E13, 1.0E15, 1.0E14, 45.0, 150.0
    rs = np.logspace(np.log10(r_in), np.log10(r_out), npoints, endpoint=True)
    nH2 = nH2_in * np.power(r_in / rs, 2.0)
    nTT = 1.0E-6 * nH2
    position = np.array([[r, 0, 0] for r in rs])
    velocity = np.array([[v_out / CC * (r - r_in) / (r_out - r_in), 0, 0] for r in rs])
    nbs, n_nbs = neighbors_1d(npoints)
    direction, weight = rays_spherical_symmetry(nrays)
    return dict(dimension=1, spherical=True, position=position, velocity=velocity,
                nbs=nbs, n_nbs=n_nbs, boundary2point=np.array([0, npoints - 1], dtype=np.uint32),
                direction=direction, weight=weight, symbols=["test", "H2", "e-"],
                abundance=np.stack([nTT, nH2, np.zeros(npoints)], axis=1),
                temperature=temp * np.ones(npoints), vturb2=(turb / CC)**2 * np.ones(npoints),
                linedata=[load_species("test")], nquads=nquads)


def spec_wind_3d(npoints=100000, nside_rays=4, nquads=11, nside_shell=None, seed=1234, spiral=False,
                 species=("co",), nlines_keep=None, v_inf=15.0e3, rho_contrast=1.0):
    """BASELINE configs 2 (100k points, nside 4, CO) and 3 (1M points, nside 8, CO + SiO-like).

    Synthetic spherically symmetric (or spiral-perturbed) wind point cloud, SURVEY 8(d):
    r = exp(U(ln 1e12, ln 1e14)) m, isotropic directions, outer HEALPix boundary shell at
    1.02e14 m, Delaunay neighbours, v = v_inf (1 - 0.9 r_in/r) r_hat, n_H2 = 1e16 (r_in/r)^2,
    n_CO = 1e-4 n_H2, T = 2500 (r_in/r)^0.6 + 10 K, v_turb = 1.5 km/s.
    """
    rng = np.random.default_rng(seed)
    r_in, r_out = 1.0e12, 1.0e14
    r = np.exp(rng.uniform(np.log(r_in), np.log(r_out), npoints))
    d = rng.normal(size=(npoints, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    if nside_shell is None:
        nside_shell = max(2, int(round(math.sqrt(max(npoints, 48) / 12.0 / 2.7) / 2)) * 2)
        nside_shell = min(nside_shell, 32)
    shell = healpix_pix2vec_ring(nside_shell) * (1.02 * r_out)
    pos_in = d * r[:, None]
    position = np.concatenate([pos_in, shell], axis=0)
    ntot = position.shape[0]
    rr = np.linalg.norm(position, axis=1)
    rhat = position / rr[:, None]
    vel = (v_inf * (1.0 - 0.9 * r_in / rr) / CC)[:, None] * rhat
    nH2 = 1.0e16 * (r_in / rr)**2
    if spiral:
        phi = np.arctan2(position[:, 1], position[:, 0])
        nH2 = nH2 * (1.0 + 0.8 * np.cos(phi - 2.0 * np.pi * rr / 2.0e13))
        tang = np.stack([-np.sin(phi), np.cos(phi), np.zeros(ntot)], axis=1)
        vel = vel + (5.0e3 * np.sqrt(r_in / rr) / CC)[:, None] * tang
    nH2 = nH2 * rho_contrast
    temperature = 2500.0 * (r_in / rr)**0.6 + 10.0
    vturb2 = (1.5e3 / CC)**2 * np.ones(ntot)
    nbs, n_nbs = neighbors_delaunay(position)
    boundary2point = np.arange(npoints, ntot, dtype=np.uint32)
    direction = antipodal_ray_set(nside_rays)
    weight = np.full(direction.shape[0], 1.0 / direction.shape[0])
    linedata, symbols, abn_cols = [], [], []
    for s in species:
        if s == "co":
            ld = load_species("co")
            x = 1.0e-4
        elif s == "sio":
            ld = scaled_species(load_species("co"), "SiO", 44.0, 21.712 / 57.636, (3.098 / 0.1098)**2)
            x = 1.0e-6
        elif s == "test":
            ld = load_species("test")
            x = 1.0e-8
        else:
            raise ValueError(s)
        if nlines_keep is not None:
            ld = restrict_lines(ld, nlines_keep)
        linedata.append(ld)
        symbols.append(ld["sym"])
        abn_cols.append(x * nH2)
    symbols = symbols + ["H2", "e-"]
    abundance = np.stack(abn_cols + [nH2, np.zeros(ntot)], axis=1)
    return dict(dimension=3, spherical=False, position=position, velocity=vel, nbs=nbs, n_nbs=n_nbs,
                boundary2point=boundary2point, direction=direction, weight=weight, symbols=symbols,
                abundance=abundance, temperature=temperature, vturb2=vturb2, linedata=linedata,
                nquads=nquads)


def with_adaptive_directions(spec, seed=99):
    """The same model with Rays::use_adaptive_directions (rays.hpp:23-31): every point gets its own rotation of the ray set (antipodal
    pairs stay pairs, as the reference assumes, rays.cpp:52-63) and its own weights (pair-symmetric, normalised per point)."""
    rng = np.random.default_rng(seed)
    out = dict(spec)
    d, w = np.asarray(spec["direction"], dtype=np.float64), np.asarray(spec["weight"], dtype=np.float64)
    N, R = len(spec["position"]), len(d)
    dirs, wts = np.empty((N, R, 3)), np.empty((N, R))
    h = R // 2
    for p in range(N):
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        dd = d @ q.T
        dd /= np.linalg.norm(dd, axis=1)[:, None]
        dd[h:] = -dd[:h]  # exact antipodes, as in the HEALPix-based sets
        f = rng.uniform(0.5, 1.5, h)
        ww = np.concatenate([w[:h] * f, w[h:] * f])
        dirs[p], wts[p] = dd, ww / ww.sum()
    if spec.get("spherical") or int(spec.get("dimension", 3)) == 1:
        raise ValueError("adaptive directions: 3D models only")
    out["direction"], out["weight"] = dirs.reshape(N * R, 3), wts.reshape(N * R)
    out["use_adaptive_directions"] = True
    return out


def build(core, io, name, spec):
    return assemble(core, io, name, **spec)


def spec_lattice_sphere(npoints=10_000_000, nside_rays=1, nquads=3, seed=1234, line_keep=(1,), v_inf=15.0e3):
    """BASELINE configs[4] (SURVEY 8d "C5 inputs"): a point cloud of ~npoints points on a jittered cubic lattice masked to a
    sphere of radius 1e14 m (spacing a, jitter U(-0.3a, 0.3a)), 26-stencil neighbour lists built in O(N) with numpy (Qhull at
    10M points is not practical), boundary = lattice points with a stencil neighbour outside the sphere, the radial laws of
    spec_wind_3d (radii below r_in = 1e12 m are clamped to r_in in the laws), CO restricted to one transition (J = 2-1)."""
    rng = np.random.default_rng(seed)
    r_in, r_out = 1.0e12, 1.0e14
    n = int(round((npoints * 6.0 / math.pi) ** (1.0 / 3.0)))          # (pi/6) n^3 lattice points inside the sphere
    a = 2.0 * r_out / n
    idx = np.arange(n, dtype=np.float64) - (n - 1) / 2.0
    inside = (idx[:, None, None] ** 2 + idx[None, :, None] ** 2 + idx[None, None, :] ** 2) <= (n / 2.0) ** 2
    ids = np.full((n + 2, n + 2, n + 2), -1, dtype=np.int64)               # one layer of padding: no wrap-around
    N = int(inside.sum())
    core = ids[1:-1, 1:-1, 1:-1]
    core[inside] = np.arange(N, dtype=np.int64)
    ii, jj, kk = np.nonzero(inside)
    position = (np.stack([idx[ii], idx[jj], idx[kk]], axis=1) + rng.uniform(-0.3, 0.3, (N, 3))) * a
    offs = [(dx, dy, dz) for dx in (-1, 0, 1) for dy in (-1, 0, 1) for dz in (-1, 0, 1) if (dx, dy, dz) != (0, 0, 0)]
    cols = np.empty((N, 26), dtype=np.int64)
    for c, (dx, dy, dz) in enumerate(offs):
        cols[:, c] = ids[ii + 1 + dx, jj + 1 + dy, kk + 1 + dz]
    valid = cols >= 0
    n_nbs = valid.sum(axis=1).astype(np.uint32)
    nbs = cols[valid].astype(np.uint32)                                   # row-major: the lists in stencil order
    boundary2point = np.nonzero(n_nbs < 26)[0].astype(np.uint32)
    del cols, valid, ids
    rr = np.maximum(np.linalg.norm(position, axis=1), r_in)
    rhat = position / np.maximum(np.linalg.norm(position, axis=1), 1e-300)[:, None]
    vel = (v_inf * (1.0 - 0.9 * r_in / rr) / CC)[:, None] * rhat
    nH2 = 1.0e16 * (r_in / rr) ** 2
    temperature = 2500.0 * (r_in / rr) ** 0.6 + 10.0
    vturb2 = (1.5e3 / CC) ** 2 * np.ones(N)
    direction = antipodal_ray_set(nside_rays)
    weight = np.full(direction.shape[0], 1.0 / direction.shape[0])
    ld = restrict_lines(load_species("co"), line_keep)
    symbols = [ld["sym"], "H2", "e-"]
    abundance = np.stack([1.0e-4 * nH2, nH2, np.zeros(N)], axis=1)
    return dict(dimension=3, spherical=False, position=position, velocity=vel, nbs=nbs, n_nbs=n_nbs, boundary2point=boundary2point,
                direction=direction, weight=weight, symbols=symbols, abundance=abundance, temperature=temperature, vturb2=vturb2,
                linedata=[ld], nquads=nquads)
