"""Python side of the host layer: build a `magritte_b200.core.Model` (C++/pybind11, magritte_b200/host/) from the plain
arrays of `flatten.flatten_model` plus the line data, or directly from any object with magritte.core's Model surface
(reference: src/bindings/pybindings.cpp:123-677). The compute_* methods then carry the reference's names:
compute_spectral_discretisation, compute_inverse_line_widths, compute_LTE_level_populations,
compute_radiation_field_feautrier_order_2, compute_Jeff, compute_level_populations_from_stateq,
compute_level_populations(use_Ng_acceleration, max_niterations), compute_image_new(rx, ry, rz, nx, ny)."""
import numpy as np

PARAMETER_NAMES = ("n_off_diag", "max_width_fraction", "convergence_fraction", "min_rel_pop_for_convergence", "pop_prec", "min_opacity",
                   "min_line_opacity", "max_interpolation_diff", "max_distance_opacity_contribution", "population_inversion_fraction", "min_dtau",
                   "use_adaptive_Ng_acceleration", "adaptive_Ng_acceleration_use_max_criterion", "Ng_acceleration_mem_limit",
                   "adaptive_Ng_acceleration_min_order", "Ng_acceleration_remove_N_its", "one_line_approximation",
                   "sum_opacity_emissivity_over_all_lines")


def load_core():
    """The compiled host module; fails loudly when magritte_b200/host/build.sh has not been run."""
    from . import core  # noqa: PLC0415
    return core


def linedata_of(model):
    """Per line-producing species: everything Lines/LineProducingSpecies/Linedata/CollisionPartner hold that the
    level-population loop reads (reference linedata.cpp:12-71, collisionPartner.cpp:10-45, quadrature)."""
    out = []
    for lsp in model.lines.lineProducingSpecies:
        ld = lsp.linedata
        d = dict(num=int(ld.num), nlev=int(ld.nlev), nrad=int(ld.nrad), inverse_mass=float(ld.inverse_mass),
                 irad=np.array(ld.irad, dtype=np.uint32), jrad=np.array(ld.jrad, dtype=np.uint32),
                 energy=np.array(ld.energy, dtype=np.float64), weight=np.array(ld.weight, dtype=np.float64),
                 frequency=np.array(ld.frequency, dtype=np.float64), A=np.array(ld.A, dtype=np.float64),
                 Ba=np.array(ld.Ba, dtype=np.float64), Bs=np.array(ld.Bs, dtype=np.float64),
                 quadrature_roots=np.array(lsp.quadrature.roots, dtype=np.float64),
                 quadrature_weights=np.array(lsp.quadrature.weights, dtype=np.float64), colpar=[])
        for c in range(int(ld.ncolpar)):
            cp = ld.colpar[c]
            d["colpar"].append(dict(num_col_partner=int(cp.num_col_partner), orth_or_para_H2=str(cp.orth_or_para_H2),
                                    icol=np.array(cp.icol, dtype=np.uint32), jcol=np.array(cp.jcol, dtype=np.uint32),
                                    tmp=np.array(cp.tmp, dtype=np.float64), Ce=np.array(cp.Ce, dtype=np.float64),
                                    Cd=np.array(cp.Cd, dtype=np.float64)))
        out.append(d)
    return out


def parameters_of(model):
    P = model.parameters
    d = {k: getattr(P, k) for k in PARAMETER_NAMES if hasattr(P, k)}
    d["spherical_symmetry"] = bool(P.spherical_symmetry())
    d["dimension"] = int(P.dimension())
    return d


def build(flat, species, abundance, parameters, device=0):
    """flat: flatten.flatten_model(..., with_state=False is enough); species: linedata_of(...); abundance [npoints][nspecs]."""
    core = load_core()
    m = core.Model()
    m.device = device
    P = m.mutable_parameters()
    for k, v in parameters.items():
        setattr(P, k, v)
    nrays = int(flat["nrays"])
    d, w = np.asarray(flat["direction"], dtype=np.float64).reshape(-1, 3), np.asarray(flat["weight"], dtype=np.float64).reshape(-1)
    m.set_geometry(flat["position"], flat["velocity"], flat["n_neighbors"], flat["neighbors"], np.ascontiguousarray(d[:nrays]), np.ascontiguousarray(w[:nrays]),
                   flat["antipod"], flat["boundary2point"], flat["boundary_condition"], flat["boundary_temperature"])
    if flat.get("use_adaptive_directions"):  # one ray set per point (Rays::use_adaptive_directions)
        m.set_adaptive_directions(d, w)
    m.set_thermodynamics(flat["temperature"], flat["vturb2"])
    m.set_abundance(np.ascontiguousarray(abundance, dtype=np.float64))
    for s in species:
        m.add_species(s)
    return m


def from_reference(model, device=0):
    """A magritte_b200.core.Model holding the same inputs as a magritte.core-style `model`."""
    from . import flatten
    flat = flatten.flatten_model(model, with_state=False)
    ab = np.array(model.chemistry.species.abundance, dtype=np.float64)
    return build(flat, linedata_of(model), ab, parameters_of(model), device)


def species_from_spec(spec):
    """The `species` list of build() from a modelbuild spec (parsed LAMDA dicts; same mapping as modelbuild.set_linedata)."""
    from .modelbuild import gauss_hermite
    roots, weights = gauss_hermite(int(spec["nquads"]))
    symbols = list(spec["symbols"])
    out = []
    for ld in spec["linedata"]:
        d = dict(num=symbols.index(ld["sym"]), nlev=int(ld["nlev"]), nrad=int(ld["nrad"]), inverse_mass=float(ld["inverse_mass"]),
                 irad=np.asarray(ld["irad"], dtype=np.uint32), jrad=np.asarray(ld["jrad"], dtype=np.uint32),
                 energy=np.asarray(ld["energy"], dtype=np.float64), weight=np.asarray(ld["weight"], dtype=np.float64),
                 frequency=np.asarray(ld["frequency"], dtype=np.float64), A=np.asarray(ld["A"], dtype=np.float64),
                 Ba=np.asarray(ld["Ba"], dtype=np.float64), Bs=np.asarray(ld["Bs"], dtype=np.float64),
                 quadrature_roots=np.asarray(roots, dtype=np.float64), quadrature_weights=np.asarray(weights, dtype=np.float64), colpar=[])
        for cp in ld["colpar"]:
            d["colpar"].append(dict(num_col_partner=symbols.index(cp["name"]) if cp["name"] in symbols else 0,
                                    orth_or_para_H2=str(cp["orth_or_para_H2"]), icol=np.asarray(cp["icol"], dtype=np.uint32),
                                    jcol=np.asarray(cp["jcol"], dtype=np.uint32), tmp=np.asarray(cp["tmp"], dtype=np.float64),
                                    Ce=np.asarray(cp["Ce"], dtype=np.float64), Cd=np.asarray(cp["Cd"], dtype=np.float64)))
        out.append(d)
    return out


def from_spec(spec, flat, device=0, **parameters):
    """Host Model of a synthetic modelbuild spec (bench.py's level-population leg); flat = synth.flat_from_spec(spec)."""
    prm = dict(spherical_symmetry=bool(spec["spherical"]), dimension=int(spec.get("dimension", 1 if spec["spherical"] else 3)))
    prm.update(parameters)
    return build(flat, species_from_spec(spec), np.asarray(spec["abundance"], dtype=np.float64), prm, device)
