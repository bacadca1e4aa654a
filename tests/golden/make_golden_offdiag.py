"""Golden Lambda operators with n_off_diag > 0 (banded ALO, Solver::update_Lambda solver.tpp:1806-1850) from the UNMODIFIED
compiled reference (oracle/_ref). Per case: the model of case_<name>.npz with parameters.n_off_diag = K, one
compute_radiation_field_feautrier_order_2 call; stored: J and the (p, line) lists (nr, Ls) in the reference's insertion order,
padded to the longest list."""
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

CASES = {"c2_vanzadelhoff": 2, "velgrad_1d": 3, "wind3d_small": 2, "c1_all_constant": 1, "wind3d_spiral_2spec": 1}


def main():
    core, mem = refcore.load()
    core.pcmt_set_n_threads_avail(8)
    specs = util.small_specs()
    for name, K in CASES.items():
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", name), name, specs[name])
            mb.prepare(m)
            if name == "wind3d_spiral_2spec":
                m.parameters.sum_opacity_emissivity_over_all_lines = False
            m.parameters.n_off_diag = K
            m.compute_radiation_field_feautrier_order_2()
        flat = flatten.flatten_model(m)
        N, L = flat["npoints"], flat["nlines"]
        lists = {}
        lsp = m.lines.lineProducingSpecies
        lid0, cap = 0, 1
        for l in range(len(lsp)):
            Ls, nr = lsp[l].Lambda.Ls, lsp[l].Lambda.nr
            for p in range(N):
                for k in range(lsp[l].linedata.nrad):
                    lists[(p, lid0 + k)] = (list(nr[p][k]), list(Ls[p][k]))
                    cap = max(cap, len(nr[p][k]))
            lid0 += lsp[l].linedata.nrad
        val = np.zeros((N, L, cap)); nrs = np.full((N, L, cap), 0xffffffff, dtype=np.uint32); cnt = np.zeros((N, L), dtype=np.uint32)
        for (p, l), (a, b) in lists.items():
            cnt[p, l] = len(a); nrs[p, l, :len(a)] = a; val[p, l, :len(b)] = b
        path = os.path.join(HERE, f"offdiag_{name}.npz")
        np.savez_compressed(path, n_off_diag=np.array(K), J=np.array(m.radiation.J), lambda_val=val, lambda_nr=nrs, lambda_cnt=cnt)
        print(name, "K", K, "cap", cap, "bytes", os.path.getsize(path))


if __name__ == "__main__":
    main()
