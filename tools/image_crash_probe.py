#!/usr/bin/env python
"""Which side of tests/test_gpu_core_compat.py::test_same_script_same_image crashes now and then? python tools/image_crash_probe.py ref|new N"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import util
from magritte_b200 import core_compat, modelbuild as mb
from oracle import refcore
which, n = sys.argv[1], int(sys.argv[2])
core = refcore.load()[0] if which == "ref" else core_compat
if which == "ref" and len(sys.argv) > 3:
    core.pcmt_set_n_threads_avail(int(sys.argv[3]))
for name in ("wind3d_small", "wind3d_spiral_2spec"):
    spec = util.small_specs()[name]
    with util.quiet():
        m = mb.build(core, core.IoPython("mem", f"probe_{name}"), f"probe_{name}", spec)
        mb.prepare(m)
    nu0 = float(np.array(m.lines.line)[0]); dv = 20.0e3 / 2.99792458e8
    for i in range(n):
        with util.quiet():
            m.compute_spectral_discretisation(nu0 * (1 - dv), nu0 * (1 + dv), 12)
            m.compute_image_new(1, 16, 12)
            m.compute_image_new(0.3, 0.5, 0.81, 16, 12)
print(which, "ok", n)
