"""ctypes front-end of the C restatement oracle (oracle/feautrier_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; never by the product package.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libfeautrier_oracle.so")


def build(force=False):
    """Compile with the reference's Release flags (reference: CMakeLists.txt:28,37): -O3, no -march, no fast-math."""
    src = os.path.join(HERE, "feautrier_oracle.c")
    if (not force) and os.path.isfile(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(src):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["gcc", "-O3", "-DNDEBUG", "-std=gnu11", "-fopenmp", "-fPIC", "-shared",
                           "-ffp-contract=off", "-fno-fast-math", src, "-o", LIB, "-lm"])
    return LIB


_u32p, _f64p, _i32p = C.POINTER(C.c_uint32), C.POINTER(C.c_double), C.POINTER(C.c_int32)


class OrcModel(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("npoints", "nrays", "hnrays", "nboundary", "nlines", "nlspecs", "nquads", "nfreqs")] + [
        ("spherical_symmetry", C.c_int32), ("approx", C.c_int32), ("n_off_diag", C.c_uint32), ("lambda_cap", C.c_uint32),
        ("position", _f64p), ("velocity", _f64p), ("n_neighbors", _u32p), ("cum_n_neighbors", _u32p), ("neighbors", _u32p),
        ("direction", _f64p), ("weight", _f64p), ("antipod", _u32p), ("point2boundary", _u32p),
        ("boundary_condition", _i32p), ("boundary_temperature", _f64p),
        ("temperature", _f64p), ("vturb2", _f64p),
        ("nrad", _u32p), ("nrad_cum", _u32p), ("inverse_mass", _f64p), ("line", _f64p), ("line_A", _f64p),
        ("quad_weights", _f64p), ("sorted_line", _f64p), ("sorted_line_map", _u32p), ("max_inverse_mass", C.c_double),
        ("nu", _f64p), ("corresponding_line", _u32p),
        ("emissivity", _f64p), ("opacity", _f64p), ("inverse_width", _f64p), ("population_tot", _f64p),
        ("min_opacity", C.c_double), ("max_distance_opacity_contribution", C.c_double),
        ("max_interpolation_diff", C.c_double), ("max_width_fraction", C.c_double),
        ("J", _f64p), ("u", _f64p), ("L_val", _f64p), ("L_nr", _u32p), ("L_cnt", _u32p),
        ("criterion", _f64p), ("dshift_max", _f64p),
        ("stat_steps", C.c_uint64), ("stat_nbs", C.c_uint64), ("stat_entries", C.c_uint64), ("max_length", C.c_uint32),
        ("adaptive_directions", C.c_int32), ("min_dtau", C.c_double)]


def _ptr(a, t):
    return a.ctypes.data_as(t)


class Oracle:
    """Owns the flat arrays (keeps them alive) and the C struct built from `flatten_model` output."""

    def __init__(self, flat, approx=None, n_off_diag=None, store_u=True, threads=None):
        from magritte_b200.flatten import approx_of
        self.lib = C.CDLL(build())
        assert self.lib.orc_sizeof_model is not None
        self.lib.orc_sizeof_model.restype = C.c_size_t
        assert self.lib.orc_sizeof_model() == C.sizeof(OrcModel)
        self.flat = f = flat
        self.threads = threads
        m = self.m = OrcModel()
        for k in ("npoints", "nrays", "hnrays", "nboundary", "nlines", "nlspecs", "nquads", "nfreqs", "spherical_symmetry"):
            setattr(m, k, int(f[k]))
        prm = f["params"]
        m.approx = approx_of(prm) if approx is None else approx
        m.n_off_diag = prm["n_off_diag"] if n_off_diag is None else n_off_diag
        m.lambda_cap = 1 + 2 * m.n_off_diag * m.hnrays if m.n_off_diag else 1
        for k, t in (("position", _f64p), ("velocity", _f64p), ("n_neighbors", _u32p), ("cum_n_neighbors", _u32p),
                     ("neighbors", _u32p), ("direction", _f64p), ("weight", _f64p), ("antipod", _u32p),
                     ("point2boundary", _u32p), ("boundary_condition", _i32p), ("boundary_temperature", _f64p),
                     ("temperature", _f64p), ("vturb2", _f64p), ("nrad", _u32p), ("nrad_cum", _u32p),
                     ("inverse_mass", _f64p), ("line", _f64p), ("line_A", _f64p), ("quad_weights", _f64p),
                     ("sorted_line", _f64p), ("sorted_line_map", _u32p), ("nu", _f64p), ("corresponding_line", _u32p),
                     ("emissivity", _f64p), ("opacity", _f64p), ("inverse_width", _f64p), ("population_tot", _f64p)):
            assert f[k].flags["C_CONTIGUOUS"], k
            setattr(m, k, _ptr(f[k], t))
        m.max_inverse_mass = f["max_inverse_mass"]
        m.adaptive_directions = 1 if f.get("use_adaptive_directions") else 0
        for k in ("min_opacity", "max_distance_opacity_contribution", "max_interpolation_diff", "max_width_fraction"):
            setattr(m, k, float(prm[k]))
        m.min_dtau = float(prm.get("min_dtau", 1.0e-15))
        N, F, L = m.npoints, m.nfreqs, m.nlines
        self.J = np.zeros((N, F))
        self.u = np.zeros((m.hnrays, N, F)) if store_u else None
        self.L_val = np.zeros((N, L, m.lambda_cap))
        self.L_nr = np.zeros((N, L, m.lambda_cap), dtype=np.uint32)
        self.L_cnt = np.zeros((N, L), dtype=np.uint32)
        self.criterion = np.zeros((N, L))
        self.dshift_max = np.zeros(N)
        m.J = _ptr(self.J, _f64p)
        m.u = _ptr(self.u, _f64p) if store_u else None
        m.L_val, m.L_nr, m.L_cnt = _ptr(self.L_val, _f64p), _ptr(self.L_nr, _u32p), _ptr(self.L_cnt, _u32p)
        m.criterion, m.dshift_max = _ptr(self.criterion, _f64p), _ptr(self.dshift_max, _f64p)
        self.lib.orc_prepare(C.byref(m))
        self.lib.orc_ray_lengths_max.restype = C.c_uint32
        self.lib.orc_trace.restype = C.c_uint32
        self._maxlen = None

    def max_length(self):
        if self._maxlen is None:
            self._maxlen = int(self.lib.orc_ray_lengths_max(C.byref(self.m)))
        return self._maxlen

    def trace(self, o, rr, l, cap=None):
        cap = cap or max(8, self.max_length())
        n = 2 * cap + 1
        nr, nri = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        fac, sh, dz = np.zeros(n), np.zeros(n), np.zeros(n)
        coff = C.c_uint32(0)
        k = self.lib.orc_trace(C.byref(self.m), C.c_uint32(o), C.c_uint32(rr), C.c_uint32(l), C.c_uint32(cap),
                               _ptr(nr, _u32p), _ptr(nri, _u32p), _ptr(fac, _f64p), _ptr(sh, _f64p), _ptr(dz, _f64p),
                               C.byref(coff))
        return dict(nr=nr[:k], nr_interp=nri[:k], interp_factor=fac[:k], shift=sh[:k], dZ=dz[:max(k - 1, 0)],
                    centre_offset=coff.value, n_tot=k)

    def anis(self):
        """Solver::solve_feautrier_order_2_anis (solver.tpp:365-456) from the stored u of a solve(): (J, J2_0, J2_1_Re, J2_1_Im, J2_2_Re,
        J2_2_Im), each [npoints][nlines], accumulated in the reference's order (rr outside z, one rounding per operation)."""
        f, m = self.flat, self.m
        assert self.u is not None, "Oracle(store_u=True) and solve() first"
        N, H, L, nq, R = m.npoints, m.hnrays, m.nlines, m.nquads, m.nrays
        d = np.asarray(f["direction"], dtype=np.float64).reshape(-1, 3)
        w = np.asarray(f["weight"], dtype=np.float64).reshape(-1)
        qw = np.repeat(np.asarray(f["quad_weights"]), np.asarray(f["nrad"]), axis=0)  # [nlines][nquads]
        sqrt3, inv_sqrt2 = np.sqrt(3.0), 1.0 / np.sqrt(2.0)
        out = np.zeros((6, N, L))
        pts = np.arange(N)
        for rr in range(H):
            idx = pts * R + rr if m.adaptive_directions else np.full(N, rr)
            nx, ny, nz, wt = d[idx, 0], d[idx, 1], d[idx, 2], w[idx]
            wts = (None, inv_sqrt2 * (3.0 * nz * nz - 1.0), -sqrt3 * nx * nz, -sqrt3 * ny * nz, 0.5 * sqrt3 * (nx * nx - ny * ny), sqrt3 * nx * ny)
            for z in range(nq):
                du = qw[None, :, z] * wt[:, None] * self.u[rr].reshape(N, L, nq)[:, :, z]
                out[0] += 2.0 * du
                for q in range(1, 6):
                    out[q] += wts[q][:, None] * du
        return out

    def solve(self, o_range=None, rr_range=None, half_length=None, twosided=False):
        """twosided=False: the reference's elimination order (the oracle). twosided=True: the product's two-sided
        re-ordering of the same solve, restated operation for operation (bit-comparable with the CUDA kernel)."""
        m = self.m
        self.lib.orc_set_twosided(C.c_int(1 if twosided else 0))
        o0, o1 = o_range or (0, m.npoints)
        r0, r1 = rr_range or (0, m.hnrays)
        hl = half_length or self.max_length()
        if self.threads:
            os.environ["OMP_NUM_THREADS"] = str(self.threads)
        rc = self.lib.orc_solve(C.byref(m), C.c_uint32(o0), C.c_uint32(o1), C.c_uint32(r0), C.c_uint32(r1), C.c_uint32(hl))
        assert rc == 0
        return self.J

    def shortchar(self, o_range=None, rr_range=None, half_length=None):
        """Solver::solve_shortchar_order_0 (solver.tpp:150-235): J, u = (I_f + I_b) / 2 and the local Lambda elements (two adds per
        (origin, ray pair, frequency), one per half ray) in the reference's order."""
        m = self.m
        o0, o1 = o_range or (0, m.npoints)
        r0, r1 = rr_range or (0, m.hnrays)
        hl = half_length or self.max_length()
        if self.threads:
            os.environ["OMP_NUM_THREADS"] = str(self.threads)
        rc = self.lib.orc_shortchar(C.byref(m), C.c_uint32(o0), C.c_uint32(o1), C.c_uint32(r0), C.c_uint32(r1), C.c_uint32(hl))
        assert rc == 0
        return self.J

    def image(self, ray_dir, nx, ny, dimension, optical_depth=False):
        """The new imager (Solver::image_feautrier_order_2_new_imager) on the model's current frequency discretisation.
        ray_dir is normed here like Model::compute_image_new does (model.cpp:1014). -> dict(ImX, ImY, I, start_pt, n_tot)."""
        m = self.m
        r = np.asarray(ray_dir, dtype=np.float64)
        r = np.ascontiguousarray(r * (1 / np.sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2])))
        npix = nx * ny
        ImX, ImY, I = np.zeros(npix), np.zeros(npix), np.zeros((npix, m.nfreqs))
        sp, nt = np.zeros(npix, np.uint32), np.zeros(npix, np.uint32)
        if self.threads:
            os.environ["OMP_NUM_THREADS"] = str(self.threads)
        fn = self.lib.orc_image_optical_depth if optical_depth else self.lib.orc_image
        rc = fn(C.byref(m), C.c_int(dimension), _ptr(r, _f64p), C.c_uint32(nx), C.c_uint32(ny), _ptr(ImX, _f64p),
                                _ptr(ImY, _f64p), _ptr(I, _f64p), _ptr(sp, _u32p), _ptr(nt, _u32p))
        assert rc == 0
        return dict(ImX=ImX, ImY=ImY, I=I, start_pt=sp, n_tot=nt)

    def set_column(self, density, o_range=None):
        """Model::set_column (model.cpp:1143-1154, Solver::get_column solver.tpp:2968-2991) -> column [nrays][o1-o0]."""
        m = self.m
        o0, o1 = o_range or (0, m.npoints)
        rho = np.ascontiguousarray(density, dtype=np.float64)
        assert rho.shape == (m.npoints,)
        col = np.zeros((m.nrays, o1 - o0))
        if self.threads:
            os.environ["OMP_NUM_THREADS"] = str(self.threads)
        rc = self.lib.orc_set_column(C.byref(m), _ptr(rho, _f64p), C.c_uint32(o0), C.c_uint32(o1), _ptr(col, _f64p))
        assert rc == 0
        return col
