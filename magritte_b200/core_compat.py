"""`magritte.core`'s Python surface over the B200 host layer: same class names, attribute tree, container behaviour
(`.set(ndarray)`, `np.array(x)`), overloads, return values and exception texts as the reference's pybind11 module
(reference src/bindings/pybindings.cpp:20-823), for the Feautrier order-2 path and its callers. A script written against
`magritte.core` -- e.g. this repository's own driver `modelbuild.build()` / `prepare()`, which is what drives the compiled
reference in the tests -- runs unchanged with `core = magritte_b200.core_compat` (tests/test_gpu_core_compat.py).

How it is layered: the classes below are the HOST data model the reference keeps in `Model` (plain numpy arrays in the
reference's layouts). `Model.read(io)` / `write(io)` use the reference's on-disk contract (the dataset and attribute paths of
src/model/**/*.cpp read()/write()) through an `Io` object; `IoPython("mem", name)` is the in-process back-end,
`IoPython("hdf5", name)` needs h5py. Every `compute_*` call syncs the tree into the C++ `magritte_b200.core.Model` (which
drives the GPU through the C ABI) and copies the results back into the tree. There is no CPU implementation here.
"""
import enum

import numpy as np

from . import flatten

CC, HH, KB, AMU, T_CMB = 2.99792458E+8, 6.62607004E-34, 1.38064852E-23, 1.66053904E-27, 2.7254800
__version__ = "0.9.3"


# ---- Paracabs multi-threading / message-passing shims (pybindings.cpp:26-42): the GPU path has no host thread pool ----
_n_threads = 1


def pcmt_n_threads_avail():
    return _n_threads


def pcmt_set_n_threads_avail(n):
    global _n_threads
    _n_threads = int(n)


def pcmp_comm_rank():
    return 0


def pcmp_comm_size():
    return 1


def pcmp_length(n):
    return int(n)


def pcmp_start(n):
    return 0


def pcmp_stop(n):
    return int(n)


class BoundaryCondition(enum.IntEnum):  # src/model/geometry/boundary/boundary.hpp
    Zero = 0
    Thermal = 1
    CMB = 2


class ImageType(enum.IntEnum):
    Intensity = 0
    OpticalDepth = 1


class ImagePointPosition(enum.IntEnum):
    AllModelPoints = 0
    ProjectionSurface = 1


# ---- containers: Paracabs Vector / Matrix / Tensor as Python sees them (pybindings.cpp:680-822) ----
class Container:
    """Buffer-protocol container with `.set(ndarray)`: np.array(x) copies out, x.set(a) copies in (dtype and rank checked)."""

    def __init__(self, dtype, ndim, owner=None, tag=None):
        self._dtype, self._ndim = np.dtype(dtype), ndim
        self._a = np.zeros((0,) * ndim, dtype=self._dtype)
        self._owner, self._tag = owner, tag

    def set(self, a):
        a = np.asarray(a)
        if a.dtype != self._dtype:
            raise TypeError(f"set(): incompatible array dtype {a.dtype}, expected {self._dtype}")  # pybind11's overload failure
        if a.ndim != self._ndim:
            raise RuntimeError("Number of dimensions must be " + str(self._ndim))
        self._a = np.array(a, dtype=self._dtype, order="C")
        if self._owner is not None:
            self._owner._touch(self._tag)

    def _assign(self, a):  # results written by the library
        self._a = np.ascontiguousarray(a, dtype=self._dtype)

    def __array__(self, dtype=None, copy=None):
        return np.array(self._a, dtype=dtype) if dtype is not None else self._a.copy()

    def __len__(self):
        return len(self._a)

    def __getitem__(self, i):
        return self._a[i]

    @property
    def shape(self):
        return self._a.shape


def VReal(owner=None, tag=None):
    return Container(np.float64, 1, owner, tag)


def VSize(owner=None, tag=None):
    return Container(np.uint32, 1, owner, tag)


def MReal(owner=None, tag=None):
    return Container(np.float64, 2, owner, tag)


def TReal(owner=None, tag=None):
    return Container(np.float64, 3, owner, tag)


def VVector3D(owner=None, tag=None):
    return Container(np.float64, 2, owner, tag)


class SetOnceError(RuntimeError):
    pass


class Parameters:
    """src/model/parameters/parameters.hpp: set-once essentials (setter / getter METHODS) and plain tunables (attributes)."""
    _ESSENTIAL = ("model_name", "dimension", "npoints", "nrays", "hnrays", "nboundary", "nfreqs", "nspecs", "nlspecs", "nlines", "nquads",
                  "use_scattering", "spherical_symmetry", "adaptive_ray_tracing")

    def __init__(self):
        self.__dict__["_once"] = {}
        d = self.__dict__
        d.update(n_off_diag=0, max_width_fraction=0.35, convergence_fraction=0.995, min_rel_pop_for_convergence=1.0e-10, pop_prec=1.0e-6,
                 min_opacity=1.0e-26, min_line_opacity=1.0e-13, min_line_emissivity=1.0e-50, min_dtau=1.0e-15, max_interpolation_diff=1.4,
                 population_inversion_fraction=-1.0, store_intensities=False, use_adaptive_Ng_acceleration=True, Ng_acceleration_mem_limit=9,
                 adaptive_Ng_acceleration_min_order=2, adaptive_Ng_acceleration_use_max_criterion=True, Ng_acceleration_remove_N_its=0,
                 one_line_approximation=False, sum_opacity_emissivity_over_all_lines=False, max_distance_opacity_contribution=10.0)
        for name in self._ESSENTIAL:
            self._make(name)

    def _make(self, name):
        def setter(value, _n=name):
            if _n in self._once and self._once[_n] != value:  # setOnce.hpp:8-16 DoubleSetException (same value is accepted)
                raise SetOnceError("Error: Trying to set a SetOnce object twice (with a different value).")
            self._once[_n] = value

        def getter(_n=name):
            if _n not in self._once:
                raise SetOnceError("Error: Trying to get a SetOnce object that was not set.")  # GetBeforeSetException
            return self._once[_n]
        self.__dict__["set_" + name] = setter
        self.__dict__[name] = getter

    def _has(self, name):
        return name in self._once


class _Node:
    """A sub-object of Model that reports changes of its containers (the device copy is refreshed before the next compute_*)."""

    def __init__(self, model):
        self._model = model

    def _touch(self, tag):
        self._model._dirty = True


class Points(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.position, self.velocity = VVector3D(self, "geometry"), VVector3D(self, "geometry")
        self.neighbors, self.n_neighbors, self.cum_n_neighbors = VSize(self, "geometry"), VSize(self, "geometry"), VSize(self, "geometry")


class Rays(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.direction, self.weight, self.antipod = VVector3D(self, "geometry"), VReal(self, "geometry"), VSize(self, "geometry")
        self.use_adaptive_directions = False


class Boundary(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.boundary2point, self.point2boundary = VSize(self, "geometry"), VSize(self, "geometry")
        self.boundary_temperature = VReal(self, "geometry")
        self._condition = []

    def set_boundary_condition(self, b, cond):
        if len(self._condition) <= b:
            self._condition.extend([BoundaryCondition.Zero] * (b + 1 - len(self._condition)))
        self._condition[b] = BoundaryCondition(cond)
        self._touch("geometry")

    def get_boundary_condition(self, b):
        return self._condition[b] if b < len(self._condition) else BoundaryCondition.Zero


class Geometry:
    def __init__(self, model):
        self.points, self.rays, self.boundary = Points(model), Rays(model), Boundary(model)


class Species(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.symbol, self._abundance = [], []

    @property
    def abundance(self):
        return self._abundance

    @abundance.setter
    def abundance(self, v):
        self._abundance = v
        self._touch("chemistry")


class Chemistry:
    def __init__(self, model):
        self.species = Species(model)


class Temperature(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.gas = VReal(self, "thermo")


class Turbulence(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.vturb2 = VReal(self, "thermo")


class Thermodynamics:
    def __init__(self, model):
        self.temperature, self.turbulence = Temperature(model), Turbulence(model)


class CollisionPartner:  # linedata/collisionPartner/collisionPartner.hpp
    def __init__(self):
        self.num_col_partner, self.orth_or_para_H2, self.ntmp, self.ncol = 0, "n", 0, 0
        self.icol, self.jcol, self.tmp, self.Ce, self.Cd = [], [], [], [], []


def vCollisionPartner(items=()):
    return list(items)


class Linedata:  # linedata/linedata.hpp
    def __init__(self):
        self.num, self.sym, self.inverse_mass, self.nlev, self.nrad = 0, "", 0.0, 0, 0
        self.irad, self.jrad, self.energy, self.weight, self.frequency, self.A, self.Ba, self.Bs = [], [], [], [], [], [], [], []
        self.ncolpar, self.colpar = 0, []


class Quadrature(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.roots, self.weights = VReal(self, "lines"), VReal(self, "lines")


class Lambda:
    """lambda/lambda.hpp: ragged lists Ls[p][k][m] / nr[p][k][m]."""

    def __init__(self):
        self.Ls, self.nr = [], []


class LineProducingSpecies:
    def __init__(self, model):
        self.linedata, self.quadrature, self.Lambda = Linedata(), Quadrature(model), Lambda()
        self.population = np.zeros(0, dtype=np.longdouble)  # VectorXld [npoints * nlev]
        self.population_tot = np.zeros(0)
        self.population_prev1 = self.population_prev2 = self.population_prev3 = np.zeros(0, dtype=np.longdouble)
        self.Jlin, self.Jeff, self.Jdif, self.J = (np.zeros((0, 0)) for _ in range(4))
        self.fraction_not_converged = self.relative_change_mean = self.relative_change_max = 0.0

    def index(self, p, i):
        return i + p * self.linedata.nlev


class Lines(_Node):
    def __init__(self, model):
        super().__init__(model)
        self.lineProducingSpecies = []
        self.line = np.zeros(0)
        self.emissivity, self.opacity, self.inverse_width = MReal(), MReal(), MReal()

    def resize_LineProducingSpecies(self, n):
        self.lineProducingSpecies = [LineProducingSpecies(self._model) for _ in range(n)]
        self._touch("lines")


class Frequencies:
    def __init__(self):
        self.nu = MReal()
        self.corresponding_line = VSize()


class Radiation:
    def __init__(self):
        self.frequencies = Frequencies()
        self.J, self.u, self.v, self.I = MReal(), TReal(), TReal(), TReal()


class Image:  # src/model/image/image.hpp as bound at pybindings.cpp:100-120
    def __init__(self, im, kind):
        self.imageType, self.imagePointPosition = kind, ImagePointPosition.ProjectionSurface
        self.nfreqs, self.freqs = int(im.nfreqs), np.asarray(im.freqs)
        self.ImX, self.ImY, self.I = np.asarray(im.ImX), np.asarray(im.ImY), np.asarray(im.I)
        self.ray_nr = -1


# ---- Io (src/io/io.hpp): the nine calls of src/io/python/io_python.cpp over a back-end ----
_MEM_FILES = {}  # IoPython("mem", name): io_file -> {"data": {path: ndarray}, "attrs": {path: value}}


class Io:
    pass


class IoPython(Io):
    """IoPython(implementation, io_file) (pybindings.cpp:70-80). "mem": process-local store with the semantics of
    src/io/python/io_python_hdf5.py (attributes addressed as "group/.name"); "hdf5": h5py files in the reference's layout."""

    def __init__(self, implementation, io_file):
        self.implementation, self.io_file = implementation, io_file
        if implementation == "hdf5":
            try:
                import h5py  # noqa: F401
            except ImportError as e:
                raise RuntimeError("IoPython('hdf5', ...) needs h5py, which is not installed; use IoPython('mem', ...)") from e
        elif implementation != "mem":
            raise RuntimeError(f"IoPython: unknown implementation '{implementation}'")

    # -- back-end primitives --
    def _store(self, create=False):
        if create:
            return _MEM_FILES.setdefault(self.io_file, {"data": {}, "attrs": {}})
        if self.io_file not in _MEM_FILES:
            raise OSError("Failed to read " + self.io_file)  # std::ios_base::failure of Model::read (model.cpp:11-16)
        return _MEM_FILES[self.io_file]

    def write_array(self, path, a):
        if self.implementation == "mem":
            self._store(True)["data"][path.strip("/")] = np.array(a)
            return
        import h5py
        with h5py.File(self.io_file, "a") as f:
            if path in f:
                del f[path]
            f.create_dataset(path, data=np.asarray(a))

    def read_array(self, path):
        if self.implementation == "mem":
            return self._store()["data"].get(path.strip("/"))
        import h5py
        with h5py.File(self.io_file, "r") as f:
            return np.array(f[path]) if path in f else None

    def write_attr(self, path, value):
        if self.implementation == "mem":
            self._store(True)["attrs"][path] = value
            return
        import h5py
        obj, attr = path.split(".")[0], path.split(".")[1]
        with h5py.File(self.io_file, "a") as f:
            (f.require_group(obj) if obj else f).attrs[attr] = value

    def read_attr(self, path, default=None):
        if self.implementation == "mem":
            return self._store()["attrs"].get(path, default)
        import h5py
        obj, attr = path.split(".")[0], path.split(".")[1]
        with h5py.File(self.io_file, "r") as f:
            g = f[obj] if obj else f
            return g.attrs.get(attr, default)

    def count(self, group, contains):
        """Number of children of `group` whose name contains `contains` (how lineProducingSpecies_ / collisionPartner_ are
        counted, io_python_hdf5.py:32-43)."""
        if self.implementation == "mem":
            st = self._store()
            pre = group.strip("/") + "/"
            kids = set()
            for key in list(st["data"].keys()) + [k.split(".")[0] for k in st["attrs"].keys()]:
                key = key.strip("/")
                if key.startswith(pre) and len(key) > len(pre):
                    kids.add(key[len(pre):].split("/")[0])
            return sum(1 for k in kids if contains in k)
        import h5py
        with h5py.File(self.io_file, "r") as f:
            return sum(1 for k in f[group].keys() if contains in k)


class IoText(Io):
    def __init__(self, io_file):
        raise RuntimeError("IoText is outside the Feautrier order-2 path of this package; use IoPython")


class Solver:  # pybindings.cpp:83-85 binds an empty class
    pass


_BOOL = {True: "true", False: "false"}
_PERSISTED_TUNABLES = ("n_off_diag", "max_width_fraction", "convergence_fraction", "min_rel_pop_for_convergence", "pop_prec", "min_opacity",
                       "Ng_acceleration_mem_limit")
_PERSISTED_BOOLS = ("store_intensities", "one_line_approximation", "use_adaptive_Ng_acceleration")


class Model:
    """magritte.core.Model (pybindings.cpp:123-271; src/model/model.hpp:19-122)."""

    def __init__(self, name=None):
        self.parameters = Parameters()
        self.geometry, self.chemistry, self.thermodynamics = Geometry(self), Chemistry(self), Thermodynamics(self)
        self.lines, self.radiation = Lines(self), Radiation()
        self.images, self.error_mean, self.error_max = [], [], []
        self.density, self.column = VReal(), MReal()
        self.device, self.device_level_populations = 0, True
        self._core, self._dirty, self._sd = None, True, None
        self._comm = None
        if name is not None:  # Model(const string name): read(IoPython("hdf5", name)) (model.hpp:35-46)
            self.parameters.set_model_name(name)
            self.read()

    # ------------------------------------------------------------------ IO: the reference's on-disk contract
    def write(self, io=None):
        P = self.parameters
        io = io or IoPython("hdf5", P.model_name())
        g, npoints = self.geometry, P.npoints()
        for k in ("dimension", "npoints", "nrays", "nboundary", "nspecs", "nlspecs", "nquads"):
            if P._has(k):
                io.write_attr("." + k, int(getattr(P, k)()))
        io.write_attr(".version", __version__)
        io.write_attr(".spherical_symmetry", _BOOL[bool(P.spherical_symmetry())])
        io.write_attr(".adaptive_ray_tracing", _BOOL[bool(g.rays.use_adaptive_directions)])
        for k in _PERSISTED_TUNABLES:
            io.write_attr("." + k, getattr(P, k))
        for k in _PERSISTED_BOOLS:
            io.write_attr("." + k, _BOOL[bool(getattr(P, k))])
        io.write_array("geometry/points/position", np.asarray(g.points.position))
        io.write_array("geometry/points/velocity", np.asarray(g.points.velocity))
        io.write_array("geometry/points/n_neighbors", np.asarray(g.points.n_neighbors))
        io.write_array("geometry/points/neighbors", np.asarray(g.points.neighbors))
        io.write_array("geometry/rays/direction", np.asarray(g.rays.direction))
        io.write_array("geometry/rays/weight", np.asarray(g.rays.weight))
        b2p = np.asarray(g.boundary.boundary2point)
        io.write_array("geometry/boundary/boundary2point", b2p)
        io.write_array("geometry/boundary/boundary_temperature", np.asarray(g.boundary.boundary_temperature))
        io.write_array("geometry/boundary/boundary_condition", np.array([int(g.boundary.get_boundary_condition(b)) for b in range(len(b2p))], dtype=np.int64))
        io.write_array("chemistry/species/abundance", np.asarray(self.chemistry.species.abundance, dtype=np.float64).reshape(npoints, -1))
        io.write_array("chemistry/species/species", np.array([str(s) for s in self.chemistry.species.symbol]))
        io.write_array("thermodynamics/temperature/gas", np.asarray(self.thermodynamics.temperature.gas))
        io.write_array("thermodynamics/turbulence/vturb2", np.asarray(self.thermodynamics.turbulence.vturb2))
        for l, ls in enumerate(self.lines.lineProducingSpecies):
            pre = f"lines/lineProducingSpecies_{l}/"
            ld = ls.linedata
            for k in ("num", "sym", "inverse_mass", "nlev", "nrad"):
                io.write_attr(pre + "linedata/." + k, getattr(ld, k))
            for k, dt in (("irad", np.int64), ("jrad", np.int64), ("energy", np.float64), ("weight", np.float64), ("frequency", np.float64),
                          ("A", np.float64), ("Bs", np.float64), ("Ba", np.float64)):
                io.write_array(pre + "linedata/" + k, np.asarray(getattr(ld, k), dtype=dt))
            for c, cp in enumerate(ld.colpar):
                cpre = pre + f"linedata/collisionPartner_{c}/"
                io.write_attr(cpre + ".num_col_partner", int(cp.num_col_partner))
                io.write_attr(cpre + ".orth_or_para_H2", str(cp.orth_or_para_H2))
                for k, dt in (("icol", np.int64), ("jcol", np.int64), ("tmp", np.float64), ("Ce", np.float64), ("Cd", np.float64)):
                    io.write_array(cpre + k, np.asarray(getattr(cp, k), dtype=dt))
            io.write_array(pre + "quadrature/roots", np.asarray(ls.quadrature.roots))
            io.write_array(pre + "quadrature/weights", np.asarray(ls.quadrature.weights))
            if len(ls.population):  # checkpoint / resume (lineProducingSpecies.cpp:138-162)
                nlev = int(ld.nlev)
                io.write_array(pre + "population", np.asarray(ls.population, dtype=np.float64).reshape(-1, nlev))
                io.write_array(pre + "population_tot", np.asarray(ls.population_tot))
                if np.asarray(ls.Jlin).size:
                    io.write_array(pre + "J_lin", np.asarray(ls.Jlin))
                    io.write_array(pre + "J_eff", np.asarray(ls.Jeff))
        if np.asarray(self.radiation.frequencies.nu).size:
            io.write_array("radiation/frequencies/nu", np.asarray(self.radiation.frequencies.nu))
        return 0

    def read(self, io=None):
        P = self.parameters
        if isinstance(io, str):
            P.set_model_name(io)
            io = None
        io = io or IoPython("hdf5", P.model_name())
        if isinstance(io, IoPython) and io.implementation == "mem":
            io._store()  # a missing "file" fails like Model::read does
        for k in ("dimension", "npoints", "nrays", "nboundary", "nspecs", "nlspecs", "nquads"):
            v = io.read_attr("." + k)
            if v is not None:
                getattr(P, "set_" + k)(int(v))
        P.set_spherical_symmetry(str(io.read_attr(".spherical_symmetry", "false")) == "true")
        for k in _PERSISTED_TUNABLES:
            v = io.read_attr("." + k)
            if v is not None:
                setattr(P, k, type(getattr(P, k))(v))
        for k in _PERSISTED_BOOLS:
            v = io.read_attr("." + k)
            if v is not None:
                setattr(P, k, str(v) == "true")
        g = self.geometry
        npoints = P.npoints()
        g.points.position.set(np.asarray(io.read_array("geometry/points/position"), dtype=np.float64).reshape(npoints, 3))
        g.points.velocity.set(np.asarray(io.read_array("geometry/points/velocity"), dtype=np.float64).reshape(npoints, 3))
        nn = np.asarray(io.read_array("geometry/points/n_neighbors"), dtype=np.uint32)
        g.points.n_neighbors.set(nn)
        g.points.neighbors.set(np.asarray(io.read_array("geometry/points/neighbors"), dtype=np.uint32))
        cum = np.zeros(npoints, dtype=np.uint64)
        cum[1:] = np.cumsum(nn[:-1], dtype=np.uint64)
        g.points.cum_n_neighbors.set(cum.astype(np.uint32))  # Points::read (points.cpp:44-60)
        d = np.asarray(io.read_array("geometry/rays/direction"), dtype=np.float64).reshape(-1, 3)
        nrays = P.nrays()
        if len(d) == nrays:  # Rays::read decides from the length of the direction array (rays.cpp:10-19)
            g.rays.use_adaptive_directions = False
        elif len(d) == nrays * npoints:
            g.rays.use_adaptive_directions = True
        else:
            P.set_nrays(len(d))  # raises like the reference: "will error, and give some more info on the sizes of the arrays"
        g.rays.direction.set(d)
        g.rays.weight.set(np.asarray(io.read_array("geometry/rays/weight"), dtype=np.float64).reshape(-1))
        P.set_hnrays(nrays // 2)
        from .synth import antipodes
        g.rays.antipod.set(antipodes(d[:nrays]))  # Rays::read (rays.cpp:52-63)
        b2p = np.asarray(io.read_array("geometry/boundary/boundary2point"), dtype=np.uint32)
        g.boundary.boundary2point.set(b2p)
        p2b = np.full(npoints, npoints, dtype=np.uint32)  # boundary.cpp:28-37: npoints marks an interior point
        p2b[b2p] = np.arange(len(b2p), dtype=np.uint32)
        g.boundary.point2boundary.set(p2b)
        g.boundary.boundary_temperature.set(np.asarray(io.read_array("geometry/boundary/boundary_temperature"), dtype=np.float64))
        bc = io.read_array("geometry/boundary/boundary_condition")
        for b in range(len(b2p)):
            g.boundary.set_boundary_condition(b, BoundaryCondition(int(bc[b])) if bc is not None else BoundaryCondition.Zero)
        self.chemistry.species.abundance = np.asarray(io.read_array("chemistry/species/abundance"), dtype=np.float64).reshape(npoints, -1).tolist()
        self.chemistry.species.symbol = [str(s) for s in np.asarray(io.read_array("chemistry/species/species")).tolist()]
        self.thermodynamics.temperature.gas.set(np.asarray(io.read_array("thermodynamics/temperature/gas"), dtype=np.float64))
        self.thermodynamics.turbulence.vturb2.set(np.asarray(io.read_array("thermodynamics/turbulence/vturb2"), dtype=np.float64))
        nl = io.count("lines", "lineProducingSpecies_")  # Lines::read (lines.cpp:11-69)
        P.set_nlspecs(nl)
        self.lines.resize_LineProducingSpecies(nl)
        line = []
        for l, ls in enumerate(self.lines.lineProducingSpecies):
            pre = f"lines/lineProducingSpecies_{l}/"
            ld = ls.linedata
            ld.num, ld.sym = int(io.read_attr(pre + "linedata/.num")), str(io.read_attr(pre + "linedata/.sym"))
            ld.inverse_mass = float(io.read_attr(pre + "linedata/.inverse_mass"))
            ld.nlev, ld.nrad = int(io.read_attr(pre + "linedata/.nlev")), int(io.read_attr(pre + "linedata/.nrad"))
            for k in ("irad", "jrad"):
                setattr(ld, k, [int(x) for x in io.read_array(pre + "linedata/" + k)])
            for k in ("energy", "weight", "frequency", "A", "Bs", "Ba"):
                setattr(ld, k, [float(x) for x in io.read_array(pre + "linedata/" + k)])
            ld.ncolpar = io.count(pre + "linedata", "collisionPartner_")
            ld.colpar = []
            for c in range(ld.ncolpar):
                cpre = pre + f"linedata/collisionPartner_{c}/"
                cp = CollisionPartner()
                cp.num_col_partner, cp.orth_or_para_H2 = int(io.read_attr(cpre + ".num_col_partner")), str(io.read_attr(cpre + ".orth_or_para_H2"))
                cp.icol, cp.jcol = [int(x) for x in io.read_array(cpre + "icol")], [int(x) for x in io.read_array(cpre + "jcol")]
                cp.tmp = [float(x) for x in io.read_array(cpre + "tmp")]
                cp.Ce, cp.Cd = np.asarray(io.read_array(cpre + "Ce")).tolist(), np.asarray(io.read_array(cpre + "Cd")).tolist()
                cp.ntmp, cp.ncol = len(cp.tmp), len(cp.icol)
                ld.colpar.append(cp)
            ls.quadrature.roots.set(np.asarray(io.read_array(pre + "quadrature/roots"), dtype=np.float64))
            ls.quadrature.weights.set(np.asarray(io.read_array(pre + "quadrature/weights"), dtype=np.float64))
            line.extend(ld.frequency)
            pop = io.read_array(pre + "population")
            if pop is not None:
                ls.population = np.asarray(pop, dtype=np.longdouble).reshape(-1)
                ls.population_tot = np.asarray(io.read_array(pre + "population_tot"), dtype=np.float64)
                self._restored = True
        self.lines.line = np.array(line, dtype=np.float64)
        P.set_nlines(len(line))
        if nl:
            P.set_nquads(len(self.lines.lineProducingSpecies[0].quadrature.roots))
        self._dirty, self._core = True, None
        return 0

    # ------------------------------------------------------------------ multi-GPU (one process per GPU; no reference counterpart)
    @staticmethod
    def nccl_unique_id():
        from . import model as hostmodel
        return hostmodel.load_core().Model.nccl_unique_id()

    def set_communicator(self, nranks, rank, uid):
        self._comm = (int(nranks), int(rank), uid)
        if self._core is not None:
            self._core.set_communicator(*self._comm)

    # ------------------------------------------------------------------ sync: host tree -> magritte_b200.core.Model -> GPU
    def _sync(self):
        from . import model as hostmodel
        if self._core is None or self._dirty:
            P = self.parameters
            if not P._has("hnrays"):
                P.set_hnrays(P.nrays() // 2)
            if not P._has("nlines"):
                P.set_nlines(sum(int(ls.linedata.nrad) for ls in self.lines.lineProducingSpecies))
            if len(np.asarray(self.geometry.rays.antipod)) != P.nrays():  # a model assembled in memory without read()
                from .synth import antipodes
                self.geometry.rays.antipod._assign(antipodes(np.asarray(self.geometry.rays.direction)[:P.nrays()]))
                b2p = np.asarray(self.geometry.boundary.boundary2point)
                p2b = np.full(P.npoints(), P.npoints(), dtype=np.uint32)
                p2b[b2p] = np.arange(len(b2p), dtype=np.uint32)
                self.geometry.boundary.point2boundary._assign(p2b)
                self.lines.line = np.array([f for ls in self.lines.lineProducingSpecies for f in ls.linedata.frequency], dtype=np.float64)
            old = self._core
            self._core = hostmodel.from_reference(self, self.device)
            if self._comm is not None:
                self._core.set_communicator(*self._comm)
            if old is not None or getattr(self, "_restored", False):  # keep what the previous core object had computed
                self._push_state()
            self._dirty = False
        c = self._core
        c.device_level_populations = bool(self.device_level_populations)
        Pc = c.mutable_parameters()
        for k in hostmodel.PARAMETER_NAMES:
            if hasattr(Pc, k):
                setattr(Pc, k, getattr(self.parameters, k))
        return c

    def _push_state(self):
        """Populations restored by read() (checkpoint / resume) or kept across a re-sync go back into the core object."""
        c = self._core
        for l, ls in enumerate(self.lines.lineProducingSpecies):
            if len(ls.population):
                pop = np.asarray(ls.population, dtype=np.longdouble)
                hi = pop.astype(np.float64)
                lo = (pop - hi.astype(np.longdouble)).astype(np.float64)
                c.set_population(l, hi, lo, np.asarray(ls.population_tot, dtype=np.float64))
        if any(len(ls.population) for ls in self.lines.lineProducingSpecies):
            c.refresh_line_state_from_populations()

    def _pull_lines(self):
        c, P = self._core, self.parameters
        self.lines.emissivity._assign(c.get_emissivity())
        self.lines.opacity._assign(c.get_opacity())
        self.lines.inverse_width._assign(c.get_inverse_width())
        for l, ls in enumerate(self.lines.lineProducingSpecies):
            hi, lo = c.get_population_hilo(l)
            ls.population = (hi.astype(np.longdouble) + lo.astype(np.longdouble)).reshape(-1)
            ls.population_tot = c.get_population_tot(l)
            ls.Jlin, ls.Jeff, ls.J = c.get_Jlin(l), c.get_Jeff(l), c.get_J_species(l)
            ls.fraction_not_converged, ls.relative_change_mean, ls.relative_change_max = c.convergence(l)

    def _pull_radiation(self, sparse=False):
        c = self._core
        if not sparse:
            self.radiation.J._assign(c.get_J())
        N = self.parameters.npoints()
        for l, ls in enumerate(self.lines.lineProducingSpecies):
            lam, cnt = c.get_lambda(l), c.get_lambda_cnt(l)
            ls.Lambda.Ls = [[[float(lam[p, k])] if cnt[p, k] else [] for k in range(lam.shape[1])] for p in range(N)]
            ls.Lambda.nr = [[[p] if cnt[p, k] else [] for k in range(lam.shape[1])] for p in range(N)]
            ls.J = c.get_J_species(l)

    # ------------------------------------------------------------------ the reference's compute_* surface (pybindings.cpp:160-262)
    def compute_inverse_line_widths(self):
        c = self._sync()
        c.compute_inverse_line_widths()
        self.lines.inverse_width._assign(c.get_inverse_width())
        return 0

    def compute_spectral_discretisation(self, *args):
        c = self._sync()
        if len(args) == 0:
            c.compute_spectral_discretisation()
            self._sd = "lines"
        elif len(args) == 1:  # (width): model.cpp:165-252, the same bins for every point around each line; not on this path
            raise RuntimeError("compute_spectral_discretisation(width) is not part of the Feautrier order-2 path of this package")
        elif len(args) == 2:
            c.compute_spectral_discretisation(float(args[0]), float(args[1]))
            self._sd = "image"
        else:
            c.compute_spectral_discretisation(float(args[0]), float(args[1]), int(args[2]))
            self._sd = "image"
        if self._sd == "lines":
            nu = c.get_nu()
            self.radiation.frequencies.nu._assign(nu)
            if not self.parameters._has("nfreqs"):
                self.parameters.set_nfreqs(int(nu.shape[1]))
        return 0

    def compute_LTE_level_populations(self):
        c = self._sync()
        c.compute_LTE_level_populations()
        self._pull_lines()
        return 0

    def compute_radiation_field_feautrier_order_2(self):
        c = self._sync()
        c.compute_radiation_field_feautrier_order_2()
        self._pull_radiation()
        return 0

    def compute_radiation_field_feautrier_order_2_sparse(self):
        c = self._sync()
        c.compute_radiation_field_feautrier_order_2_sparse()
        self._pull_radiation(sparse=True)
        return 0

    def compute_radiation_field_feautrier_order_2_anis(self):
        c = self._sync()
        c.compute_radiation_field_feautrier_order_2_anis()
        for l, ls in enumerate(self.lines.lineProducingSpecies):
            ls.J, ls.J2_0, ls.J2_1_Re, ls.J2_1_Im, ls.J2_2_Re, ls.J2_2_Im = c.get_J_anis(l)
        return 0

    def compute_radiation_field_shortchar_order_0(self):
        c = self._sync()
        c.compute_radiation_field_shortchar_order_0()
        self._pull_radiation()
        return 0

    def compute_Jeff(self):
        c = self._sync()
        c.compute_Jeff()
        self._pull_lines()
        return 0

    def compute_Jeff_sparse(self):
        c = self._sync()
        c.compute_Jeff_sparse()
        self._pull_lines()
        return 0

    def compute_level_populations_from_stateq(self):
        c = self._sync()
        c.compute_level_populations_from_stateq()
        self._pull_lines()
        return 0

    def compute_level_populations(self, use_Ng_acceleration, max_niterations):
        c = self._sync()
        n = c.compute_level_populations(bool(use_Ng_acceleration), int(max_niterations))
        self._pull_lines()
        self._pull_radiation()
        self.error_mean, self.error_max = list(c.error_mean), list(c.error_max)
        return n

    def compute_level_populations_sparse(self, use_Ng_acceleration, max_niterations):
        c = self._sync()
        n = c.compute_level_populations_sparse(bool(use_Ng_acceleration), int(max_niterations))
        self._pull_lines()
        self._pull_radiation(sparse=True)
        self.error_mean, self.error_max = list(c.error_mean), list(c.error_max)
        return n

    def compute_level_populations_shortchar(self, use_Ng_acceleration, max_niterations):
        c = self._sync()
        n = c.compute_level_populations_shortchar(bool(use_Ng_acceleration), int(max_niterations))
        self._pull_lines()
        self._pull_radiation()
        self.error_mean, self.error_max = list(c.error_mean), list(c.error_max)
        return n

    def _ray_of(self, args):
        if len(args) == 3:  # (ray_nr, Nxpix, Nypix): the direction of ray ray_nr (model.cpp:985-997)
            d = np.asarray(self.geometry.rays.direction)[int(args[0])]
            return float(d[0]), float(d[1]), float(d[2]), int(args[1]), int(args[2])
        if len(args) == 5:
            return float(args[0]), float(args[1]), float(args[2]), int(args[3]), int(args[4])
        raise TypeError("compute_image_new(ray_nr, Nxpix, Nypix) or compute_image_new(rx, ry, rz, Nxpix, Nypix)")

    def compute_image_new(self, *args):
        if len(args) == 1:  # (ray_nr): 256 x 256 pixels (model.cpp:976-983)
            args = (args[0], 256, 256)
        c = self._sync()
        c.compute_image_new(*self._ray_of(args))
        self.images.append(Image(c.images[-1], ImageType.Intensity))
        return 0

    def compute_image_optical_depth_new(self, *args):
        if len(args) == 1:
            args = (args[0], 256, 256)
        c = self._sync()
        c.compute_image_optical_depth_new(*self._ray_of(args))
        self.images.append(Image(c.images[-1], ImageType.OpticalDepth))
        return 0

    def set_column(self):
        c = self._sync()
        c.density = np.asarray(self.density)
        c.set_column()
        self.column._assign(c.column)
        return 0

    # the reference binds these too; they are other solvers (SURVEY.md 2.1 "non-target solvers")
    def _not_on_path(self, *a, **k):
        raise RuntimeError("this solver is outside the Feautrier order-2 path magritte_b200 implements")
    compute_radiation_field_feautrier_order_2_uv = compute_image = _not_on_path
    compute_image_optical_depth = compute_image_for_point = set_eta_and_chi = set_boundary_condition = _not_on_path
