"""ctypes binding of libmagritte_b200.so (the C ABI declared in include/magritte_b200.h).

This is the Python-side stub a maintainer of the reference would add next to
`magritte.core`; the library itself has no Python or torch dependency.
There is no CPU fallback: loading fails loudly if the shared library has not been built, and
`Context()` raises if there is no CUDA device.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MGB_LIB") or os.path.join(_HERE, "lib", "libmagritte_b200.so")  # MGB_LIB: A/B builds of the same library

EXPORTS = [
    "mgb_abi_version", "mgb_create", "mgb_destroy", "mgb_last_error", "mgb_set_geometry", "mgb_set_adaptive_directions", "mgb_set_thermodynamics",
    "mgb_set_lines", "mgb_set_frequencies", "mgb_set_line_state", "mgb_params_default", "mgb_set_params",
    "mgb_run_feautrier_order_2", "mgb_run_shortchar_order_0", "mgb_solve_pipeline_info", "mgb_get_J", "mgb_get_lambda", "mgb_get_u", "mgb_lambda_capacity", "mgb_get_lambda_lists", "mgb_solve_feautrier_order_2",
    "mgb_result_ptrs_dev", "mgb_invalidate_paths", "mgb_timer_start", "mgb_timer_stop", "mgb_nccl_unique_id", "mgb_comm_init", "mgb_partition", "mgb_partition_balanced", "mgb_partition_rows",
    "mgb_allgather_results", "mgb_comm_destroy", "mgb_comm_agree", "mgb_set_radiation_field", "mgb_lp_set_parameters", "mgb_lp_set_species", "mgb_lp_add_collision_partner", "mgb_lp_set_abundance",
    "mgb_lp_set_populations", "mgb_lp_get_populations", "mgb_lp_refresh_line_state", "mgb_lp_get_line_state", "mgb_lp_compute_Jeff", "mgb_lp_get_Jeff",
    "mgb_lp_update_statistical_equilibrium", "mgb_lp_ng_gram", "mgb_lp_ng_combine", "mgb_lp_finish_update", "mgb_lp_trial_convergence",
    "mgb_image_feautrier_order_2", "mgb_trace_dump", "mgb_selftest_libm", "mgb_measure_fp64_peak", "mgb_selftest_ddiv", "mgb_get_J_lines", "mgb_get_J_anis", "mgb_image_optical_depth", "mgb_set_column", "mgb_set_column_noalloc",
]


class Params(C.Structure):
    _fields_ = [("min_opacity", C.c_double), ("max_distance_opacity_contribution", C.c_double),
                ("max_interpolation_diff", C.c_double), ("max_width_fraction", C.c_double),
                ("n_off_diag", C.c_uint32), ("store_intensities", C.c_uint32),
                ("path_cache_bytes", C.c_uint64), ("scratch_bytes", C.c_uint64),
                ("cache_paths", C.c_uint32), ("anisotropic", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("n_steps", C.c_uint64), ("n_nbs_scanned", C.c_uint64), ("n_entries", C.c_uint64),
                ("n_solves", C.c_uint64), ("n_large_shift", C.c_uint64), ("max_ray_length", C.c_uint32),
                ("n_chunks", C.c_uint32), ("n_launches", C.c_uint32), ("paths_cached", C.c_uint32),
                ("ms_trace_count", C.c_float), ("ms_trace_fill", C.c_float), ("ms_solve", C.c_float),
                ("ms_reduce", C.c_float), ("ms_total", C.c_float), ("n_solve_launches", C.c_uint32),
                ("reserved", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


class ImageStats(C.Structure):
    _fields_ = [("n_pixels", C.c_uint64), ("n_entries", C.c_uint64), ("n_solves", C.c_uint64), ("max_ray_length", C.c_uint32),
                ("start_bdy", C.c_uint32), ("n_chunks", C.c_uint32), ("n_launches", C.c_uint32), ("ms_trace", C.c_float),
                ("ms_solve", C.c_float), ("ms_total", C.c_float), ("reserved", C.c_uint32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "reserved"}


_lib = None


def load():
    """Load the shared library (fails loudly when it has not been built: no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built: run magritte_b200/csrc/build.sh (or __graft_entry__.build())")
        _lib = C.CDLL(LIB_PATH)
        _lib.mgb_last_error.restype = C.c_char_p
        _lib.mgb_last_error.argtypes = [C.c_void_p]
        _lib.mgb_destroy.restype = None
        _lib.mgb_destroy.argtypes = [C.c_void_p]
    return _lib


def partition_rows(npoints, nranks, rank):
    """(o_begin, o_end, rows_per_rank) of mgb_partition_rows: host arithmetic, needs no GPU."""
    a, b, per = C.c_uint32(), C.c_uint32(), C.c_uint32()
    if load().mgb_partition_rows(C.c_uint32(npoints), C.c_int(nranks), C.c_int(rank), C.byref(a), C.byref(b), C.byref(per)) != 0:
        raise ValueError("mgb_partition_rows: invalid arguments")
    return a.value, b.value, per.value


class MgbError(RuntimeError):
    pass


def _p(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a, a.ctypes.data_as(C.c_void_p)


class Context:
    """Owns one mgb_ctx. All arrays are host numpy arrays in the reference's layouts."""

    def __init__(self, device=0):
        self.lib = load()
        self._ctx = C.c_void_p()
        rc = self.lib.mgb_create(C.c_int(device), C.byref(self._ctx))
        if rc != 0:
            raise MgbError(f"mgb_create failed ({rc}): {self.lib.mgb_last_error(None).decode()}")
        self.params = Params()
        self.lib.mgb_params_default(C.byref(self.params))
        self.npoints = self.nfreqs = self.nlines = self.hnrays = 0
        self._keep = {}

    def close(self):
        if self._ctx:
            self.lib.mgb_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise MgbError(f"{what} failed ({rc}): {self.lib.mgb_last_error(self._ctx).decode()}")

    # ---- upload -----------------------------------------------------------------------------
    def set_params(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.params, k):
                raise KeyError(k)
            setattr(self.params, k, v)
        self._check(self.lib.mgb_set_params(self._ctx, C.byref(self.params)), "mgb_set_params")

    def set_model(self, flat, frequencies=True):
        """Upload everything from a `flatten.flatten_model` dict. frequencies=False skips the line discretisation (an imaging
        model whose frequencies are passed to image() instead)."""
        f = flat
        self.npoints, self.nlines, self.nfreqs = int(f["npoints"]), int(f["nlines"]), int(f["nfreqs"])
        self.hnrays, self.nrays = int(f["hnrays"]), int(f["nrays"])
        prm = f["params"]
        self.set_params(min_opacity=prm["min_opacity"],
                        max_distance_opacity_contribution=prm["max_distance_opacity_contribution"],
                        max_interpolation_diff=prm["max_interpolation_diff"], max_width_fraction=prm["max_width_fraction"],
                        n_off_diag=int(prm["n_off_diag"]))
        pos, ppos = _p(f["position"], np.float64)
        vel, pvel = _p(f["velocity"], np.float64)
        nn, pnn = _p(f["n_neighbors"], np.uint32)
        nb, pnb = _p(f["neighbors"], np.uint32)
        adaptive = bool(f.get("use_adaptive_directions", False))  # direction / weight then hold one ray set per point
        d, pd = _p(np.asarray(f["direction"]).reshape(-1, 3)[:self.nrays], np.float64)
        w, pw = _p(np.asarray(f["weight"]).reshape(-1)[:self.nrays], np.float64)
        ap, pap = _p(f["antipod"], np.uint32)
        p2b, pp2b = _p(f["point2boundary"], np.uint32)
        bc, pbc = _p(f["boundary_condition"], np.int32)
        bt, pbt = _p(f["boundary_temperature"], np.float64)
        self._check(self.lib.mgb_set_geometry(self._ctx, C.c_uint32(self.npoints), ppos, pvel, pnn, pnb,
                                              C.c_uint64(int(nn.sum(dtype=np.uint64))), C.c_uint32(self.nrays), pd, pw, pap,
                                              C.c_uint32(int(f["nboundary"])), pp2b, pbc, pbt,
                                              C.c_int(int(f["spherical_symmetry"]))), "mgb_set_geometry")
        if adaptive:
            da, pda = _p(f["direction"], np.float64)
            wa, pwa = _p(f["weight"], np.float64)
            assert da.size == 3 * self.npoints * self.nrays and wa.size == self.npoints * self.nrays
            self._check(self.lib.mgb_set_adaptive_directions(self._ctx, pda, pwa), "mgb_set_adaptive_directions")
        T, pT = _p(f["temperature"], np.float64)
        vt, pvt = _p(f["vturb2"], np.float64)
        self._check(self.lib.mgb_set_thermodynamics(self._ctx, pT, pvt), "mgb_set_thermodynamics")
        nrad, pnrad = _p(f["nrad"], np.uint32)
        im, pim = _p(f["inverse_mass"], np.float64)
        ln, pln = _p(f["line"], np.float64)
        A, pA = _p(f["line_A"], np.float64)
        qw, pqw = _p(f["quad_weights"], np.float64)
        self._check(self.lib.mgb_set_lines(self._ctx, C.c_uint32(int(f["nlspecs"])), pnrad, pim, pln, pA,
                                           C.c_uint32(int(f["nquads"])), pqw), "mgb_set_lines")
        if frequencies:
            nu, pnu = _p(f["nu"], np.float64)
            cl, pcl = _p(f["corresponding_line"], np.uint32)
            self._check(self.lib.mgb_set_frequencies(self._ctx, C.c_uint32(self.nfreqs), pnu, pcl), "mgb_set_frequencies")
        self.set_line_state(f["emissivity"], f["opacity"], f["inverse_width"], f["population_tot"])

    def set_line_state(self, emissivity, opacity, inverse_width, population_tot):
        e, pe = _p(emissivity, np.float64)
        o, po = _p(opacity, np.float64)
        w, pw = _p(inverse_width, np.float64)
        t, pt = _p(population_tot, np.float64)
        self._check(self.lib.mgb_set_line_state(self._ctx, pe, po, pw, pt), "mgb_set_line_state")

    # ---- run --------------------------------------------------------------------------------
    def set_column(self, density, o_range=None, with_time=False):
        """mgb_set_column: Model::set_column for origins o_range -> column [nrays][o1-o0] (and the kernel's device ms)."""
        rho, prho = _p(density, np.float64)
        o0, o1 = o_range or (0, self.npoints)
        col = np.empty((self.nrays, max(o1 - o0, 0)))
        ms = C.c_float(0)
        self._check(self.lib.mgb_set_column(self._ctx, prho, C.c_uint32(o0), C.c_uint32(o1), col.ctypes.data_as(C.c_void_p),
                                            C.byref(ms)), "mgb_set_column")
        return (col, ms.value) if with_time else col

    def set_column_noalloc(self, density, o_range=None):
        """mgb_set_column_noalloc: set_column staged through the library's static device array (no device allocation in the call)."""
        rho, prho = _p(density, np.float64)
        o0, o1 = o_range or (0, self.npoints)
        col = np.empty((self.nrays, max(o1 - o0, 0)))
        self._check(self.lib.mgb_set_column_noalloc(self._ctx, prho, C.c_uint32(o0), C.c_uint32(o1), col.ctypes.data_as(C.c_void_p)),
                    "mgb_set_column_noalloc")
        return col

    def image(self, approx, dimension, ray_dir, nx, ny, nu, pix_range=None, optical_depth=False):
        """mgb_image_feautrier_order_2 (or mgb_image_optical_depth): nu is [nfreqs] (common bins) or [npoints][nfreqs].
        -> (ImX, ImY, I[pixels][nfreqs], stats)"""
        fn = self.lib.mgb_image_optical_depth if optical_depth else self.lib.mgb_image_feautrier_order_2
        nu, pnu = _p(nu, np.float64)
        per_point = 1 if nu.ndim == 2 else 0
        nfreqs = nu.shape[-1]
        d, pd = _p(ray_dir, np.float64)
        p0, p1 = pix_range or (0, nx * ny)
        X, Y, I = np.empty(nx * ny), np.empty(nx * ny), np.empty((p1 - p0, nfreqs))
        st = ImageStats()
        self._check(fn(self._ctx, C.c_int(approx), C.c_int(dimension), pd, C.c_uint32(nx), C.c_uint32(ny),
                                                         C.c_uint32(nfreqs), pnu, C.c_int(per_point), C.c_uint64(p0), C.c_uint64(p1),
                                                         X.ctypes.data_as(C.c_void_p), Y.ctypes.data_as(C.c_void_p),
                                                         I.ctypes.data_as(C.c_void_p), C.byref(st)), "mgb_image_feautrier_order_2")
        return X, Y, I, st

    def run(self, approx=2, o_range=None):
        o0, o1 = o_range or (0, self.npoints)
        st = Stats()
        self._check(self.lib.mgb_run_feautrier_order_2(self._ctx, C.c_int(approx), C.c_uint32(o0), C.c_uint32(o1),
                                                       C.byref(st)), "mgb_run_feautrier_order_2")
        self.last_range = (o0, o1)
        return st

    def pipeline_info(self):
        """-> dict(kernel, ms_records, ms_line_integrals, ms_sweep, n_ranges) of the last run (mgb_solve_pipeline_info)."""
        k, a, b, c, n = C.c_int(0), C.c_float(0), C.c_float(0), C.c_float(0), C.c_uint32(0)
        self._check(self.lib.mgb_solve_pipeline_info(self._ctx, C.byref(k), C.byref(a), C.byref(b), C.byref(c), C.byref(n)), "mgb_solve_pipeline_info")
        return dict(kernel=k.value, ms_records=a.value, ms_line_integrals=b.value, ms_sweep=c.value, n_ranges=n.value)

    def run_shortchar(self, approx=2, o_range=None, min_dtau=1.0e-15, store_u=None):
        """Solver::solve_shortchar_order_0 (solver.tpp:150-235) -> Stats; results through get_J / get_u / get_lambda_diag."""
        o0, o1 = o_range or (0, self.npoints)
        if store_u is not None:
            self.set_params(store_intensities=1 if store_u else 0)
        st = Stats()
        self._check(self.lib.mgb_run_shortchar_order_0(self._ctx, C.c_int(approx), C.c_double(min_dtau), C.c_uint32(o0), C.c_uint32(o1),
                                                       C.byref(st)), "mgb_run_shortchar_order_0")
        self.last_range = (o0, o1)
        return st

    def get_lambda_lists(self, o_range=None):
        """n_off_diag > 0: (val [rows][nlines][cap], nr, cnt) -- the reference's (p, line) lists in insertion order."""
        o0, o1 = o_range or self.last_range
        cap = C.c_uint32()
        self._check(self.lib.mgb_lambda_capacity(self._ctx, C.byref(cap)), "mgb_lambda_capacity")
        val = np.empty((o1 - o0, self.nlines, cap.value))
        nr = np.empty((o1 - o0, self.nlines, cap.value), dtype=np.uint32)
        cnt = np.empty((o1 - o0, self.nlines), dtype=np.uint32)
        self._check(self.lib.mgb_get_lambda_lists(self._ctx, C.c_uint32(o0), C.c_uint32(o1), val.ctypes.data_as(C.c_void_p),
                                                  nr.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p)), "mgb_get_lambda_lists")
        return val, nr, cnt

    def get_J(self, o_range=None):
        o0, o1 = o_range or self.last_range
        J = np.empty((o1 - o0, self.nfreqs))
        self._check(self.lib.mgb_get_J(self._ctx, C.c_uint32(o0), C.c_uint32(o1), J.ctypes.data_as(C.c_void_p)), "mgb_get_J")
        return J

    def get_J_lines(self, o_range=None):
        """lspec.J of the memory-sparse solver: J integrated over each line's quadrature, [rows][nlines]."""
        o0, o1 = o_range or self.last_range
        J = np.empty((o1 - o0, self.nlines))
        self._check(self.lib.mgb_get_J_lines(self._ctx, C.c_uint32(o0), C.c_uint32(o1), J.ctypes.data_as(C.c_void_p)), "mgb_get_J_lines")
        return J

    def get_J_anis(self, o_range=None):
        """(J, J2_0, J2_1_Re, J2_1_Im, J2_2_Re, J2_2_Im) per line, each [rows][nlines] (run with set_params(anisotropic=1))."""
        o0, o1 = o_range or self.last_range
        out = np.empty((6, o1 - o0, self.nlines))
        self._check(self.lib.mgb_get_J_anis(self._ctx, C.c_uint32(o0), C.c_uint32(o1), out.ctypes.data_as(C.c_void_p)), "mgb_get_J_anis")
        return out

    def get_lambda(self, o_range=None):
        o0, o1 = o_range or self.last_range
        val = np.empty((o1 - o0, self.nlines))
        cnt = np.empty((o1 - o0, self.nlines), dtype=np.uint32)
        self._check(self.lib.mgb_get_lambda(self._ctx, C.c_uint32(o0), C.c_uint32(o1), val.ctypes.data_as(C.c_void_p),
                                            cnt.ctypes.data_as(C.c_void_p)), "mgb_get_lambda")
        return val, cnt

    def get_u(self):
        o0, o1 = self.last_range
        u = np.empty((self.hnrays, o1 - o0, self.nfreqs))
        self._check(self.lib.mgb_get_u(self._ctx, C.c_uint32(o0), C.c_uint32(o1), u.ctypes.data_as(C.c_void_p)), "mgb_get_u")
        return u

    def solve(self, emissivity, opacity, inverse_width, population_tot, approx=2, o_range=None, J=None, lam=None):
        """mgb_solve_feautrier_order_2: host buffers in, host buffers out (the end-to-end call)."""
        o0, o1 = o_range or (0, self.npoints)
        e, pe = _p(emissivity, np.float64)
        o, po = _p(opacity, np.float64)
        w, pw = _p(inverse_width, np.float64)
        t, pt = _p(population_tot, np.float64)
        if J is None:
            J = np.empty((o1 - o0, self.nfreqs))
        if lam is None:
            lam = (np.empty((o1 - o0, self.nlines)), np.empty((o1 - o0, self.nlines), dtype=np.uint32))
        st = Stats()
        self._check(self.lib.mgb_solve_feautrier_order_2(
            self._ctx, C.c_int(approx), pe, po, pw, pt, C.c_uint32(o0), C.c_uint32(o1), J.ctypes.data_as(C.c_void_p),
            lam[0].ctypes.data_as(C.c_void_p), lam[1].ctypes.data_as(C.c_void_p), None, C.byref(st)),
            "mgb_solve_feautrier_order_2")
        self.last_range = (o0, o1)
        return J, lam, st

    def invalidate_paths(self):
        self._check(self.lib.mgb_invalidate_paths(self._ctx), "mgb_invalidate_paths")

    def timer_start(self):
        self._check(self.lib.mgb_timer_start(self._ctx), "mgb_timer_start")

    def timer_stop(self):
        ms = C.c_float()
        self._check(self.lib.mgb_timer_stop(self._ctx, C.byref(ms)), "mgb_timer_stop")
        return ms.value

    # ---- multi-GPU --------------------------------------------------------------------------
    def nccl_unique_id(self):
        buf = (C.c_char * 128)()
        rc = self.lib.mgb_nccl_unique_id(buf)
        if rc != 0:
            raise MgbError("mgb_nccl_unique_id failed")
        return bytes(buf)

    def comm_init(self, nranks, rank, uid):
        self._check(self.lib.mgb_comm_init(self._ctx, C.c_int(nranks), C.c_int(rank), C.c_char_p(uid)), "mgb_comm_init")

    def partition(self, nranks, rank, balanced=False):
        a, b = C.c_uint32(), C.c_uint32()
        fn = self.lib.mgb_partition_balanced if balanced else self.lib.mgb_partition
        self._check(fn(self._ctx, C.c_int(nranks), C.c_int(rank), C.byref(a), C.byref(b)), "mgb_partition")
        return a.value, b.value

    def allgather(self):
        self._check(self.lib.mgb_allgather_results(self._ctx), "mgb_allgather_results")
        self.last_range = (0, self.npoints)

    # ---- level populations on the device (mgb_lp_*) ----------------------------------------------
    def lp_setup(self, species, abundance, parameters=None):
        """species: the list model.linedata_of / species_from_spec builds; abundance [npoints][nspecs]."""
        prm = parameters or {}
        self._check(self.lib.mgb_lp_set_parameters(self._ctx, C.c_double(float(prm.get("min_line_opacity", 1e-13))),
                                                   C.c_double(float(prm.get("min_line_emissivity", 1e-50))),
                                                   C.c_double(float(prm.get("population_inversion_fraction", -1.0))),
                                                   C.c_double(float(prm.get("min_rel_pop_for_convergence", 1e-10)))), "mgb_lp_set_parameters")
        self.lp_nlev, self.lp_nrad = [], []
        for l, d in enumerate(species):
            a = {k: _p(d[k], np.uint32 if k in ("irad", "jrad") else np.float64) for k in ("irad", "jrad", "energy", "weight", "frequency", "A", "Ba", "Bs")}
            self._check(self.lib.mgb_lp_set_species(self._ctx, C.c_uint32(l), C.c_uint32(int(d["num"])), C.c_uint32(int(d["nlev"])), C.c_uint32(int(d["nrad"])),
                                                    a["irad"][1], a["jrad"][1], a["energy"][1], a["weight"][1], a["frequency"][1], a["A"][1], a["Ba"][1],
                                                    a["Bs"][1]), "mgb_lp_set_species")
            self.lp_nlev.append(int(d["nlev"])); self.lp_nrad.append(int(d["nrad"]))
            for cp in d["colpar"]:
                ic, pic = _p(cp["icol"], np.uint32); jc, pjc = _p(cp["jcol"], np.uint32)
                tm, ptm = _p(cp["tmp"], np.float64); Ce, pCe = _p(cp["Ce"], np.float64); Cd, pCd = _p(cp["Cd"], np.float64)
                oop = {"n": 0, "o": 1, "p": 2}[str(cp["orth_or_para_H2"])]
                self._check(self.lib.mgb_lp_add_collision_partner(self._ctx, C.c_uint32(l), C.c_uint32(int(cp["num_col_partner"])), C.c_int(oop),
                                                                  C.c_uint32(tm.size), C.c_uint32(ic.size), pic, pjc, ptm, pCe, pCd), "mgb_lp_add_collision_partner")
        ab, pab = _p(abundance, np.float64)
        self._check(self.lib.mgb_lp_set_abundance(self._ctx, C.c_uint32(ab.shape[1]), pab), "mgb_lp_set_abundance")

    def lp_set_populations(self, l, population, population_tot, lo=None):
        hi, phi = _p(population, np.float64)
        plo = None
        if lo is not None:
            lo, plo = _p(lo, np.float64)
        t, pt = _p(population_tot, np.float64)
        self._check(self.lib.mgb_lp_set_populations(self._ctx, C.c_uint32(l), phi, plo, pt), "mgb_lp_set_populations")

    def lp_get_populations(self, l):
        hi = np.empty((self.npoints, self.lp_nlev[l])); lo = np.empty_like(hi)
        self._check(self.lib.mgb_lp_get_populations(self._ctx, C.c_uint32(l), hi.ctypes.data_as(C.c_void_p), lo.ctypes.data_as(C.c_void_p)), "mgb_lp_get_populations")
        return hi, lo

    def lp_refresh_line_state(self):
        self._check(self.lib.mgb_lp_refresh_line_state(self._ctx), "mgb_lp_refresh_line_state")

    def lp_get_line_state(self):
        e, o, w = (np.empty((self.npoints, self.nlines)) for _ in range(3))
        self._check(self.lib.mgb_lp_get_line_state(self._ctx, *(a.ctypes.data_as(C.c_void_p) for a in (e, o, w))), "mgb_lp_get_line_state")
        return e, o, w

    def set_radiation_field(self, J, lam_val, lam_cnt, J_lines=None):
        pj = pjl = None
        if J is not None:
            J, pj = _p(J, np.float64)
        if J_lines is not None:
            J_lines, pjl = _p(J_lines, np.float64)
        lv, plv = _p(lam_val, np.float64)
        lc, plc = _p(lam_cnt, np.uint32)
        self._check(self.lib.mgb_set_radiation_field(self._ctx, pj, pjl, plv, plc), "mgb_set_radiation_field")
        self.last_range = (0, self.npoints)

    def lp_compute_Jeff(self, sparse=False):
        self._check(self.lib.mgb_lp_compute_Jeff(self._ctx, C.c_int(1 if sparse else 0)), "mgb_lp_compute_Jeff")

    def lp_get_Jeff(self, l):
        a, b = np.empty((self.npoints, self.lp_nrad[l])), np.empty((self.npoints, self.lp_nrad[l]))
        self._check(self.lib.mgb_lp_get_Jeff(self._ctx, C.c_uint32(l), a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p)), "mgb_lp_get_Jeff")
        return a, b

    def lp_update_statistical_equilibrium(self, pop_prec=1e-6, keep_history=False):
        """-> ([(fraction_not_converged, relative_change_mean, relative_change_max) per species], device ms)"""
        conv = (C.c_double * (3 * len(self.lp_nlev)))()
        ms = C.c_float()
        self._check(self.lib.mgb_lp_update_statistical_equilibrium(self._ctx, C.c_int(1 if keep_history else 0), C.c_double(pop_prec), conv, C.byref(ms)),
                    "mgb_lp_update_statistical_equilibrium")
        return [tuple(conv[3 * l:3 * l + 3]) for l in range(len(self.lp_nlev))], ms.value

    # ---- test hooks -------------------------------------------------------------------------
    def trace(self, o, rr, l, cap=4096):
        nr, nri = np.zeros(cap, np.uint32), np.zeros(cap, np.uint32)
        fac, sh, dz = np.zeros(cap), np.zeros(cap), np.zeros(cap)
        n, c = C.c_uint32(), C.c_uint32()
        self._check(self.lib.mgb_trace_dump(self._ctx, C.c_uint32(o), C.c_uint32(rr), C.c_uint32(l), C.c_uint32(cap),
                                            nr.ctypes.data_as(C.c_void_p), nri.ctypes.data_as(C.c_void_p),
                                            fac.ctypes.data_as(C.c_void_p), sh.ctypes.data_as(C.c_void_p),
                                            dz.ctypes.data_as(C.c_void_p), C.byref(n), C.byref(c)), "mgb_trace_dump")
        k = n.value
        return dict(nr=nr[:k], nr_interp=nri[:k], interp_factor=fac[:k], shift=sh[:k], dZ=dz[:max(k - 1, 0)],
                    centre_offset=c.value, n_tot=k)

    def libm(self, fn, x, y=None):
        x, px = _p(x, np.float32)
        out = np.empty_like(x)
        py = None
        if y is not None:
            y, py = _p(y, np.float32)
        code = {"expf": 0, "logf": 1, "erff": 2, "expm1f": 3, "powf": 4}[fn]
        self._check(self.lib.mgb_selftest_libm(self._ctx, C.c_int(code), C.c_uint64(x.size), px, py,
                                               out.ctypes.data_as(C.c_void_p)), "mgb_selftest_libm")
        return out

    def ddiv(self, a, b):
        a, pa = _p(a, np.float64)
        b, pb = _p(b, np.float64)
        out = np.empty_like(a)
        self._check(self.lib.mgb_selftest_ddiv(self._ctx, C.c_uint64(a.size), pa, pb, out.ctypes.data_as(C.c_void_p)), "mgb_selftest_ddiv")
        return out

    def fp64_peak_tflops(self):
        v = C.c_double()
        self._check(self.lib.mgb_measure_fp64_peak(self._ctx, C.byref(v)), "mgb_measure_fp64_peak")
        return v.value
