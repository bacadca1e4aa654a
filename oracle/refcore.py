"""Loader for the compiled, unmodified reference (oracle/_ref/magritte/core.so).

TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline). Nothing in the product
package imports this module.
"""
import importlib
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
_REF = os.path.join(HERE, "_ref")
_IO = os.path.join(HERE, "ref_home", "src", "io", "python")


def available():
    return os.path.isfile(os.path.join(_REF, "magritte", "core.so"))


def load():
    """Return (core module, io_python_mem module)."""
    if not available():
        raise RuntimeError("oracle/_ref/magritte/core.so missing: run oracle/build_ref.sh")
    for p in (_IO, _REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    mem = importlib.import_module("io_python_mem")
    core = importlib.import_module("magritte.core")
    return core, mem


def load_hooks():
    for p in (_IO, _REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    load()
    return importlib.import_module("magritte.ref_hooks")


class TraceHook:
    """numpy-facing wrapper of ref_hooks.TraceHook (arrays travel as raw bytes)."""

    def __init__(self, model):
        import numpy as np
        self._np = np
        self._h = load_hooks().TraceHook(model)
        self.length = self._h.length()
        self.centre = self._h.centre()

    def trace(self, o, rr, l):
        np = self._np
        d = dict(self._h.trace(int(o), int(rr), int(l)))
        for k in ("nr", "nr_interp"):
            d[k] = np.frombuffer(d[k], dtype=np.uint32).copy()
        for k in ("interp_factor", "shift", "dZ"):
            d[k] = np.frombuffer(d[k], dtype=np.float64).copy()
        return d


def dshift_max(model):
    import numpy as np
    return np.frombuffer(load_hooks().dshift_max(model), dtype=np.float64).copy()
