"""Where an image call's wall time goes: the bench model (wind, 11 quadrature points), images before and after a Feautrier pass on a
few origins (the pass sizes the big solve buffers), with and without a busy host. python tools/image_diag.py [--torch]"""
import json, sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    if "--torch" in sys.argv:
        import torch
        torch.zeros(1, device="cuda")
    from magritte_b200 import cabi, modelbuild as mb, synth
    flat = synth.flat_from_spec(mb.spec_wind_3d(npoints=100000, nside_rays=4, nquads=11, seed=1234))
    nu0 = float(flat["line"][1]); dv = 20.0e3 / 2.99792458e8
    nu = np.linspace(nu0 * (1 - dv), nu0 * (1 + dv), 100)
    ctx = cabi.Context(0)
    ctx.set_model(flat)

    def images(tag):
        for d in [(0.3, 0.5, 0.81), (1.0, 0.0, 0.0), (0.3, 0.5, 0.81)]:
            t0 = time.perf_counter()
            _, _, I, st = ctx.image(2, 3, d, 512, 512, nu)
            wall = (time.perf_counter() - t0) * 1e3
            print(json.dumps({"tag": tag, "wall_ms": round(wall, 2), "total": round(st.ms_total, 2), "trace": round(st.ms_trace, 2),
                              "solve": round(st.ms_solve, 2), "launches": st.n_launches, "chunks": st.n_chunks}))
    images("fresh")
    st = ctx.run(2, (0, 4000))
    print(json.dumps({"run_ms": st.ms_total}))
    images("after a solve")
    ctx.close()


if __name__ == "__main__":
    main()
