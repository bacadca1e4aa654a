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
