"""Shared helpers of the test-suite (test infrastructure only)."""
import contextlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


@contextlib.contextmanager
def quiet():
    """Silence the C++ side's std::cout chatter (the reference prints per ray, solver.tpp:477)."""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    try:
        os.dup2(devnull, 1)
        yield
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


def small_specs():
    """The small parity cases: BASELINE configs[0], configs[1], a Doppler-shifted 1D model and a
    small 3D Delaunay wind (the shape of configs[2], shrunk so the CPU sides finish in seconds)."""
    from magritte_b200 import modelbuild as mb
    return {
        "c1_all_constant": mb.spec_all_constant_1d(),
        "c2_vanzadelhoff": mb.spec_vanzadelhoff_1a_1d(),
        "velgrad_1d": mb.spec_velocity_gradient_1d(),
        "wind3d_small": mb.spec_wind_3d(npoints=1500, nside_rays=1, nquads=5, nlines_keep=range(6)),
        "wind3d_spiral_2spec": mb.spec_wind_3d(npoints=900, nside_rays=1, nquads=3, spiral=True, species=("co", "sio"),
                                              nlines_keep=range(4), v_inf=60.0e3),
    }


def relerr(a, b, floor=1e-300):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    d = np.abs(a - b)
    s = np.maximum(np.abs(b), floor)
    return float((d / s).max()) if d.size else 0.0


def lambda_dense(model, nlines):
    """Reference Lambda lists -> dict[(p, lid)] = sorted [(nr, Ls)]. reference: lambda.hpp:14-22"""
    out = {}
    lsp = model.lines.lineProducingSpecies
    lid0 = 0
    for l in range(len(lsp)):
        Ls, nr = lsp[l].Lambda.Ls, lsp[l].Lambda.nr
        nrad = lsp[l].linedata.nrad
        for p in range(len(Ls)):
            for k in range(nrad):
                out[(p, lid0 + k)] = sorted(zip(list(nr[p][k]), list(Ls[p][k])))
        lid0 += nrad
    return out


GOLDEN_CASES = ["c1_all_constant", "c2_vanzadelhoff", "velgrad_1d", "wind3d_small", "wind3d_spiral_2spec"]


def load_golden(name):
    """-> (flat inputs dict as magritte_b200.flatten.flatten_model builds it, dict of reference outputs)."""
    z = np.load(os.path.join(GOLDEN, f"case_{name}.npz"))
    flat, ref, params = {}, {}, {}
    for k in z.files:
        v = z[k]
        if k.startswith("param_"):
            params[k[6:]] = v.item()
        elif k.startswith("ref_") or k.startswith("trace_"):
            ref[k] = v
        else:
            flat[k] = v.item() if v.ndim == 0 else np.ascontiguousarray(v)
    flat["params"] = params
    return flat, ref


def golden_traces(ref):
    """Yield ((o, rr, l), dict of arrays) for the ray paths stored in a golden file."""
    pos = {k: 0 for k in ("nr", "nr_interp", "interp_factor", "shift", "dZ")}
    for (o, rr, l), (n, coff) in zip(ref["trace_sel"], ref["trace_meta"]):
        t = {}
        for k in pos:
            cnt = int(n) - 1 if k == "dZ" else int(n)
            cnt = max(cnt, 0)
            t[k] = ref["trace_" + k][pos[k]:pos[k] + cnt]
            pos[k] += cnt
        t["n_tot"], t["centre_offset"] = int(n), int(coff)
        yield (int(o), int(rr), int(l)), t


def assert_trace_equal(got, want, ctx=""):
    assert got["n_tot"] == want["n_tot"], f"{ctx}: n_tot {got['n_tot']} != {want['n_tot']}"
    assert got["centre_offset"] == want["centre_offset"], ctx
    for k in ("nr", "nr_interp"):
        assert np.array_equal(got[k], want[k]), f"{ctx}: {k} differs"
    for k in ("interp_factor", "shift", "dZ"):
        assert np.array_equal(np.asarray(got[k]).view(np.uint64), np.asarray(want[k]).view(np.uint64)), f"{ctx}: {k} not bit-identical"


def load_golden_image(name):
    """-> (flat inputs with the image frequency discretisation, dict(dimension, nx, ny, directions, images=[(ImX, ImY, I)]))
    from tests/golden/image_<name>.npz (tests/golden/make_golden_image.py; the model is the one of case_<name>.npz)."""
    flat, _ = load_golden(name)
    z = np.load(os.path.join(GOLDEN, f"image_{name}.npz"))
    nu = np.ascontiguousarray(z["nu"])
    flat["nfreqs"] = int(nu.size)
    flat["nu"] = np.ascontiguousarray(np.broadcast_to(nu, (flat["npoints"], nu.size)))
    flat["nu_image"] = nu
    flat["corresponding_line"] = np.zeros(nu.size, dtype=np.uint32)
    dirs = z["directions"]
    img = dict(dimension=int(z["dimension"]), nx=int(z["nx"]), ny=int(z["ny"]), directions=dirs,
               images=[(z[f"ImX_{i}"], z[f"ImY_{i}"], z[f"I_{i}"]) for i in range(len(dirs))])
    return flat, img


def load_golden_image_tau(name):
    """Optical-depth images of the same model, bins, directions and pixels as load_golden_image (tests/golden/make_golden_image_tau.py):
    -> [(ImX, ImY, tau)] per direction."""
    z = np.load(os.path.join(GOLDEN, f"imagetau_{name}.npz"))
    n = len([k for k in z.files if k.startswith("tau_")])
    return [(z[f"ImX_{i}"], z[f"ImY_{i}"], z[f"tau_{i}"]) for i in range(n)]


def load_golden_column(name):
    """-> (density [npoints], column [nrays][npoints]) of Model::set_column on the model of case_<name>.npz
    (tests/golden/make_golden_column.py)."""
    z = np.load(os.path.join(GOLDEN, f"column_{name}.npz"))
    return np.ascontiguousarray(z["density"]), np.ascontiguousarray(z["column"])


def load_golden_offdiag(name):
    """-> (flat inputs with params.n_off_diag = K, dict(J, lambda_val/nr/cnt [N][L][cap] in the reference's insertion order))."""
    flat, _ = load_golden(name)
    z = np.load(os.path.join(GOLDEN, f"offdiag_{name}.npz"))
    flat["params"] = dict(flat["params"], n_off_diag=int(z["n_off_diag"]))
    return flat, {k: z[k] for k in ("J", "lambda_val", "lambda_nr", "lambda_cnt")}


def lambda_lists_close(val, nr, cnt, ref, tol):
    """Compare (p, line) lists after sorting by nr (SURVEY.md 8c protocol); tol = 0 demands identical bits in insertion order."""
    assert np.array_equal(cnt, ref["lambda_cnt"]), "list lengths differ"
    N, L = cnt.shape
    cap = ref["lambda_nr"].shape[2]
    worst = 0.0
    for p in range(N):
        for l in range(L):
            c = int(cnt[p, l])
            a_nr, a_v = nr[p, l, :c], val[p, l, :c]
            b_nr, b_v = ref["lambda_nr"][p, l, :c], ref["lambda_val"][p, l, :c]
            if tol == 0:
                assert np.array_equal(a_nr, b_nr) and np.array_equal(a_v, b_v), (p, l)
                continue
            ia, ib = np.argsort(a_nr, kind="stable"), np.argsort(b_nr, kind="stable")
            assert np.array_equal(a_nr[ia], b_nr[ib]), (p, l)
            worst = max(worst, relerr(a_v[ia], b_v[ib]))
    assert worst <= tol, worst
    return worst
