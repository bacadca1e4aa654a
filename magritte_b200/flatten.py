"""Flatten a `magritte.core`-style Model into the plain arrays the C ABI takes.

Works on any object exposing the reference's pybind surface (reference:
src/bindings/pybindings.cpp:123-677): used on this package's own `core.Model`, and by the
tests on the compiled reference to feed identical inputs to both sides.
"""
import numpy as np


def heapsort_lines(line):
    """Sorted line list + index map, as the reference builds them in Lines::read
    (reference: src/model/lines/lines.cpp:52-64 with src/tools/heapsort.hpp). Line centre
    frequencies are distinct in practice; a stable argsort then yields the same order."""
    idx = np.argsort(line, kind="stable").astype(np.uint32)
    return np.ascontiguousarray(line[idx]), idx


def flatten_model(model, with_state=True):
    P = model.parameters
    npoints, nrays = P.npoints(), P.nrays()
    out = dict(npoints=npoints, nrays=nrays, hnrays=nrays // 2, nboundary=P.nboundary(),
               nlspecs=P.nlspecs(), nlines=P.nlines(), nquads=P.nquads(),
               spherical_symmetry=int(P.spherical_symmetry()))
    g = model.geometry
    out["position"] = np.ascontiguousarray(np.array(g.points.position), dtype=np.float64).reshape(npoints, 3)
    out["velocity"] = np.ascontiguousarray(np.array(g.points.velocity), dtype=np.float64).reshape(npoints, 3)
    out["n_neighbors"] = np.ascontiguousarray(np.array(g.points.n_neighbors), dtype=np.uint32)
    out["neighbors"] = np.ascontiguousarray(np.array(g.points.neighbors), dtype=np.uint32)
    cum = np.zeros(npoints, dtype=np.uint64)
    cum[1:] = np.cumsum(out["n_neighbors"][:-1], dtype=np.uint64)
    out["cum_n_neighbors"] = cum.astype(np.uint32)
    out["direction"] = np.ascontiguousarray(np.array(g.rays.direction), dtype=np.float64).reshape(-1, 3)
    out["weight"] = np.ascontiguousarray(np.array(g.rays.weight), dtype=np.float64)
    out["antipod"] = np.ascontiguousarray(np.array(g.rays.antipod), dtype=np.uint32)
    out["use_adaptive_directions"] = bool(g.rays.use_adaptive_directions)
    b2p = np.ascontiguousarray(np.array(g.boundary.boundary2point), dtype=np.uint32)
    out["boundary2point"] = b2p
    out["point2boundary"] = np.ascontiguousarray(np.array(g.boundary.point2boundary), dtype=np.uint32)
    out["boundary_condition"] = np.array(
        [int(g.boundary.get_boundary_condition(b)) for b in range(len(b2p))], dtype=np.int32)
    out["boundary_temperature"] = np.ascontiguousarray(np.array(g.boundary.boundary_temperature), dtype=np.float64)
    out["temperature"] = np.ascontiguousarray(np.array(model.thermodynamics.temperature.gas), dtype=np.float64)
    out["vturb2"] = np.ascontiguousarray(np.array(model.thermodynamics.turbulence.vturb2), dtype=np.float64)
    lsp = model.lines.lineProducingSpecies
    nl = P.nlspecs()
    out["nrad"] = np.array([lsp[l].linedata.nrad for l in range(nl)], dtype=np.uint32)
    out["nrad_cum"] = np.concatenate([[0], np.cumsum(out["nrad"])[:-1]]).astype(np.uint32)
    out["inverse_mass"] = np.array([lsp[l].linedata.inverse_mass for l in range(nl)], dtype=np.float64)
    out["max_inverse_mass"] = float(max(0.0, out["inverse_mass"].max()))
    out["line"] = np.ascontiguousarray(np.array(model.lines.line), dtype=np.float64)
    out["line_A"] = np.concatenate([np.array(lsp[l].linedata.A, dtype=np.float64) for l in range(nl)])
    out["quad_weights"] = np.stack([np.array(lsp[l].quadrature.weights, dtype=np.float64) for l in range(nl)])
    out["quad_roots"] = np.stack([np.array(lsp[l].quadrature.roots, dtype=np.float64) for l in range(nl)])
    out["sorted_line"], out["sorted_line_map"] = heapsort_lines(out["line"])
    p = P
    out["params"] = dict(min_opacity=p.min_opacity, max_distance_opacity_contribution=p.max_distance_opacity_contribution,
                         max_interpolation_diff=p.max_interpolation_diff, max_width_fraction=p.max_width_fraction,
                         n_off_diag=int(p.n_off_diag), one_line_approximation=bool(p.one_line_approximation),
                         sum_opacity_emissivity_over_all_lines=bool(p.sum_opacity_emissivity_over_all_lines))
    if with_state:
        nfreqs = P.nfreqs()
        out["nfreqs"] = nfreqs
        out["nu"] = np.ascontiguousarray(np.array(model.radiation.frequencies.nu), dtype=np.float64).reshape(npoints, nfreqs)
        nquads = P.nquads()
        if nfreqs == P.nlines() * nquads:
            out["corresponding_line"] = (np.arange(nfreqs, dtype=np.uint32) // nquads).astype(np.uint32)
        else:
            out["corresponding_line"] = np.zeros(nfreqs, dtype=np.uint32)
        out["emissivity"] = np.ascontiguousarray(np.array(model.lines.emissivity), dtype=np.float64)
        out["opacity"] = np.ascontiguousarray(np.array(model.lines.opacity), dtype=np.float64)
        out["inverse_width"] = np.ascontiguousarray(np.array(model.lines.inverse_width), dtype=np.float64)
        out["population_tot"] = np.stack([np.array(lsp[l].population_tot, dtype=np.float64) for l in range(nl)])
    return out


def approx_of(params):
    """ApproximationType dispatch of reference src/model/model.cpp:366-389: 0 None, 1 OneLine, 2 CloseLines."""
    if params["one_line_approximation"]:
        return 1
    if params["sum_opacity_emissivity_over_all_lines"]:
        return 0
    return 2
