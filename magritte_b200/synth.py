"""Flat hot-path inputs straight from a model spec (numpy), without a `Model` object.

Given one of the `modelbuild.spec_*` dictionaries this produces exactly the dictionary that
`flatten.flatten_model` extracts from a prepared reference `Model` (read -> compute_spectral_discretisation ->
compute_inverse_line_widths -> compute_LTE_level_populations), i.e. the arguments of the C ABI. The arithmetic
follows the reference's host code, including its long-double level populations:
  * LTE populations         src/model/lines/lineProducingSpecies/lineProducingSpecies.tpp:52-72
  * emissivity / opacity    lineProducingSpecies.tpp:23-44, src/model/lines/lines.tpp:29-41
  * inverse line widths     lines.tpp:45-59, src/model/thermodynamics/thermodynamics.tpp:27-35
  * frequencies nu(p, f)    src/model/model.cpp:70-161
  * rays antipodes          src/model/geometry/rays/rays.cpp:52-63
  * boundary maps           src/model/geometry/boundary/boundary.cpp:5-48
tests/test_synth_vs_reference.py pins it bit-for-bit against the compiled reference.
"""
import math

import numpy as np

from .lamda import AMU, HH, KB

CC_SQUARED = 8.98755179E+16
TWO_KB_OVER_AMU_CC_SQUARED = 2.0 * KB / (AMU * CC_SQUARED)
FOUR_PI = 12.566370614359172  # constants.hpp: 4*pi
HH_OVER_FOUR_PI = HH / FOUR_PI
DEFAULT_PARAMS = dict(min_opacity=1.0e-26, max_distance_opacity_contribution=10.0, max_interpolation_diff=1.4,
                      max_width_fraction=0.35, n_off_diag=0, one_line_approximation=False,
                      sum_opacity_emissivity_over_all_lines=False)


def antipodes(direction):
    """rays.cpp:52-63: last r2 with |d1 + d2|^2 < 1e-9 wins."""
    d = np.asarray(direction)
    n = len(d)
    ap = np.zeros(n, dtype=np.uint32)
    for r1 in range(n):
        s = d + d[r1]
        hit = np.nonzero((s * s).sum(axis=1) < 1.0e-9)[0]
        if len(hit):
            ap[r1] = hit[-1]
    return ap


def lte_populations(ld, abundance_col, temperature):
    """update_using_LTE; returns population [npoints][nlev] as long double (the reference's Eigen VectorXld)."""
    w = np.asarray(ld["weight"], dtype=np.float64)
    E = np.asarray(ld["energy"], dtype=np.float64)
    T = np.asarray(temperature, dtype=np.float64)
    arg = -E[None, :] / (KB * T[:, None])
    # libm exp() element by element: numpy's vectorised exp is not bit-identical to glibc's
    ex = np.fromiter(map(math.exp, arg.ravel()), dtype=np.float64, count=arg.size).reshape(arg.shape)
    pop = (w[None, :] * ex).astype(np.longdouble)  # double expression, stored long double
    part = np.zeros(len(T), dtype=np.float64)
    for i in range(len(w)):  # Real partition_function += (long double): add in long double, round to double each step
        part = (part.astype(np.longdouble) + pop[:, i]).astype(np.float64)
    tot = np.asarray(abundance_col, dtype=np.float64)
    pop = pop * (tot / part)[:, None].astype(np.longdouble)
    return pop, tot


def line_state(ld, pop, min_line_opacity=1.0e-13, min_line_emissivity=1.0e-50):
    freq = np.asarray(ld["frequency"], dtype=np.float64)
    A, Ba, Bs = (np.asarray(ld[k], dtype=np.float64) for k in ("A", "Ba", "Bs"))
    irad, jrad = np.asarray(ld["irad"], dtype=np.int64), np.asarray(ld["jrad"], dtype=np.int64)
    c = (freq * HH_OVER_FOUR_PI)  # double * double
    emis = (c * A)[None, :].astype(np.longdouble) * pop[:, irad]
    emis = np.maximum(emis, np.longdouble(min_line_emissivity)).astype(np.float64)
    opac = c[None, :].astype(np.longdouble) * (pop[:, jrad] * Ba[None, :].astype(np.longdouble)
                                               - pop[:, irad] * Bs[None, :].astype(np.longdouble))
    opac = np.maximum(opac, np.longdouble(min_line_opacity)).astype(np.float64)
    return emis, opac


def flat_from_spec(spec, params=None, min_line_opacity=1.0e-13, with_state=True):
    """with_state=False: geometry, thermodynamics and the static line tables only (what magritte_b200.model.build needs; the host
    Model computes nu, the inverse widths, the LTE populations and the line state itself)."""
    position = np.ascontiguousarray(spec["position"], dtype=np.float64)
    npoints = len(position)
    direction = np.ascontiguousarray(spec["direction"], dtype=np.float64)
    adaptive = bool(spec.get("use_adaptive_directions", False))
    nrays = len(direction) // (npoints if adaptive else 1)  # adaptive: [npoints * nrays][3] (rays.cpp:13-19)
    b2p = np.asarray(spec["boundary2point"], dtype=np.uint32)
    p2b = np.full(npoints, npoints, dtype=np.uint32)
    p2b[b2p] = np.arange(len(b2p), dtype=np.uint32)
    bc_name = spec.get("boundary_condition", "CMB")
    bt = spec.get("boundary_temperature")
    n_nbs = np.asarray(spec["n_nbs"], dtype=np.uint32)
    cum = np.zeros(npoints, dtype=np.uint64)
    cum[1:] = np.cumsum(n_nbs[:-1], dtype=np.uint64)
    T = np.ascontiguousarray(spec["temperature"], dtype=np.float64)
    vt2 = np.ascontiguousarray(spec["vturb2"], dtype=np.float64)
    linedata = spec["linedata"]
    nquads = int(spec["nquads"])
    from .modelbuild import gauss_hermite
    roots, weights = gauss_hermite(nquads)
    symbols = list(spec["symbols"])
    abundance = np.asarray(spec["abundance"], dtype=np.float64)
    nrad = np.array([int(ld["nrad"]) for ld in linedata], dtype=np.uint32)
    nlines = int(nrad.sum())
    out = dict(npoints=npoints, nrays=nrays, hnrays=nrays // 2, nboundary=len(b2p), nlspecs=len(linedata), nlines=nlines,
               nquads=nquads, nfreqs=nlines * nquads, spherical_symmetry=int(bool(spec["spherical"])),
               position=position, velocity=np.ascontiguousarray(spec["velocity"], dtype=np.float64),
               n_neighbors=n_nbs, neighbors=np.ascontiguousarray(spec["nbs"], dtype=np.uint32),
               cum_n_neighbors=cum.astype(np.uint32), direction=direction,
               weight=np.ascontiguousarray(spec["weight"], dtype=np.float64), antipod=antipodes(direction[:nrays]),
               use_adaptive_directions=adaptive, boundary2point=b2p, point2boundary=p2b,
               boundary_condition=np.full(len(b2p), {"Zero": 0, "Thermal": 1, "CMB": 2}[bc_name], dtype=np.int32),
               boundary_temperature=np.full(len(b2p), 2.72548) if bt is None else np.asarray(bt, dtype=np.float64),
               temperature=T, vturb2=vt2, nrad=nrad,
               nrad_cum=np.concatenate([[0], np.cumsum(nrad)[:-1]]).astype(np.uint32),
               inverse_mass=np.array([float(ld["inverse_mass"]) for ld in linedata]),
               line=np.concatenate([np.asarray(ld["frequency"], dtype=np.float64) for ld in linedata]),
               line_A=np.concatenate([np.asarray(ld["A"], dtype=np.float64) for ld in linedata]),
               quad_weights=np.stack([weights] * len(linedata)), quad_roots=np.stack([roots] * len(linedata)))
    out["max_inverse_mass"] = float(max(0.0, out["inverse_mass"].max()))
    from .flatten import heapsort_lines
    out["sorted_line"], out["sorted_line_map"] = heapsort_lines(out["line"])
    prm = dict(DEFAULT_PARAMS)
    if params:
        prm.update(params)
    out["params"] = prm
    if not with_state:
        return out
    emis, opac, invw, nu, ptot, pops = [], [], [], [], [], []
    for l, ld in enumerate(linedata):
        num = symbols.index(ld["sym"])
        pop, tot = lte_populations(ld, abundance[:, num], T)
        e, o = line_state(ld, pop, min_line_opacity=min_line_opacity)
        freq = np.asarray(ld["frequency"], dtype=np.float64)
        pw = np.sqrt(TWO_KB_OVER_AMU_CC_SQUARED * float(ld["inverse_mass"]) * T + vt2)  # profile_width(inverse_mass, p)
        width = freq[None, :] * pw[:, None]
        invw.append(1.0 / width)
        nu.append((freq[None, :, None] + width[:, :, None] * roots[None, None, :]).reshape(npoints, -1))
        emis.append(e)
        opac.append(o)
        ptot.append(tot)
        pops.append(pop)
    out["emissivity"] = np.ascontiguousarray(np.concatenate(emis, axis=1))
    out["opacity"] = np.ascontiguousarray(np.concatenate(opac, axis=1))
    out["inverse_width"] = np.ascontiguousarray(np.concatenate(invw, axis=1))
    out["nu"] = np.ascontiguousarray(np.concatenate(nu, axis=1))
    out["population_tot"] = np.ascontiguousarray(np.stack(ptot))
    out["corresponding_line"] = (np.arange(out["nfreqs"], dtype=np.uint32) // nquads).astype(np.uint32)
    out["_populations"] = pops
    return out
