"""Generate the golden input/output vectors of the Feautrier order-2 path from the UNMODIFIED compiled
reference (oracle/_ref, built by oracle/build_ref.sh from /root/reference). Run in the build container;
the .npz files are committed, the reference is not needed to use them.

Per case the file holds every flat input array of the path (magritte_b200.flatten.flatten_model) and the
reference's outputs: radiation.J, radiation.u (small cases), the Lambda operator lists, Model.dshift_max and
a sample of ray paths (nr_, nr_interp_, interp_factor_, shift_, dZ_) dumped through oracle/ref_hooks.cpp.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import util  # noqa: E402
from magritte_b200 import flatten, modelbuild as mb  # noqa: E402
from oracle import refcore  # noqa: E402


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    rng = np.random.default_rng(7)
    for name, spec in util.small_specs().items():
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", name), name, spec)
            mb.prepare(m)
            if name == "wind3d_spiral_2spec":
                m.parameters.sum_opacity_emissivity_over_all_lines = False
            m.compute_radiation_field_feautrier_order_2()
        flat = flatten.flatten_model(m)
        out = {}
        for k, v in flat.items():
            if k == "params":
                for pk, pv in v.items():
                    out["param_" + pk] = np.array(pv)
            else:
                out[k] = np.asarray(v)
        J = np.array(m.radiation.J)
        out["ref_J"] = J
        u = np.array(m.radiation.u)
        if u.size <= 400000:
            out["ref_u"] = u
        lam = util.lambda_dense(m, flat["nlines"])
        val = np.zeros((flat["npoints"], flat["nlines"]))
        cnt = np.zeros((flat["npoints"], flat["nlines"]), dtype=np.uint32)
        for (p, l), lst in lam.items():
            cnt[p, l] = len(lst)
            if lst:
                assert lst[0][0] == p and len(lst) == 1
                val[p, l] = lst[0][1]
        out["ref_lambda_val"], out["ref_lambda_cnt"] = val, cnt
        with util.quiet():
            out["ref_dshift_max"] = refcore.dshift_max(m)
            hook = refcore.TraceHook(m)
        out["ref_length"] = np.array(hook.length)
        N, H, L = flat["npoints"], flat["hnrays"], flat["nlines"]
        ntr = min(40, N * H * L)
        sel = np.stack([rng.integers(0, N, ntr), rng.integers(0, H, ntr), rng.integers(0, L, ntr)], axis=1).astype(np.uint32)
        sel[0] = (0, 0, 0)
        sel[-1] = (N - 1, H - 1, L - 1)
        out["trace_sel"] = sel
        cat = {k: [] for k in ("nr", "nr_interp", "interp_factor", "shift", "dZ")}
        meta = []
        for (o, rr, l) in sel:
            with util.quiet():
                t = hook.trace(o, rr, l)
            meta.append((len(t["nr"]), t["centre"] - t["first"]))
            for k in cat:
                cat[k].append(t[k])
        out["trace_meta"] = np.array(meta, dtype=np.uint32)
        for k in cat:
            out["trace_" + k] = np.concatenate(cat[k])
        path = os.path.join(HERE, f"case_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "J", J.shape, "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
