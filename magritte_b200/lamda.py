"""LAMDA line-data file parser (host-side model construction helper).

Plays the role of `magritte.setup.set_linedata_from_LAMDA_file`
(reference: magritte/setup.py:632-776), which relies on astroquery -- absent from this
image -- so the format is parsed directly. Unit conversions follow the reference:
levels re-based to 0, energies cm^-1 -> J (x 1e2*h*c), collisional rates cm^3 -> m^3
(x 1e-6), Ce = Cd * g_i/g_j * exp(-(E_i-E_j)/(k T)), nu = (E_i-E_j)/h,
Bs = A c^2 / (2 h nu^3), Ba = g_i/g_j * Bs.
"""
import re
import numpy as np

CC = 2.99792458E+8
HH = 6.62607004E-34
KB = 1.38064852E-23
AMU = 1.66053904E-27
T_CMB = 2.7254800


def parse_lamda(path):
    """Parse a LAMDA file into a plain dict of numpy arrays (SI units, 0-based levels)."""
    with open(path) as fh:
        lines = [ln.rstrip("\n") for ln in fh]
    it = iter(lines)

    def nxt():
        ln = next(it)
        while ln.strip().startswith("!"):
            ln = next(it)
        return ln

    sym = nxt().split()[0]
    mass = float(nxt().split()[0])
    nlev = int(nxt().split()[0])
    energy = np.zeros(nlev)
    weight = np.zeros(nlev)
    for i in range(nlev):
        t = nxt().split()
        energy[i] = float(t[1])
        weight[i] = float(t[2])
    nrad = int(nxt().split()[0])
    irad = np.zeros(nrad, dtype=np.uint32)
    jrad = np.zeros(nrad, dtype=np.uint32)
    A = np.zeros(nrad)
    for k in range(nrad):
        t = nxt().split()
        irad[k] = int(t[1]) - 1
        jrad[k] = int(t[2]) - 1
        A[k] = float(t[3])
    energy = energy * (1.0E+2 * HH * CC)
    ncolpar = int(nxt().split()[0])
    colpar = []
    for _ in range(ncolpar):
        desc = nxt()
        pid = int(desc.split()[0])
        # LAMDA partner codes: 1 H2, 2 p-H2, 3 o-H2, 4 e, 5 H, 6 He, 7 H+
        name = {1: "H2", 2: "H2", 3: "H2", 4: "e-", 5: "H", 6: "He", 7: "H+"}.get(pid, "H2")
        orth_or_para = {2: "p", 3: "o"}.get(pid, "n")
        ncol = int(nxt().split()[0])
        ntmp = int(nxt().split()[0])
        tmp = np.array([float(x) for x in nxt().split()])
        assert len(tmp) == ntmp
        icol = np.zeros(ncol, dtype=np.uint32)
        jcol = np.zeros(ncol, dtype=np.uint32)
        Cd = np.zeros((ntmp, ncol))
        for k in range(ncol):
            t = nxt().split()
            icol[k] = int(t[1]) - 1
            jcol[k] = int(t[2]) - 1
            Cd[:, k] = [float(x) for x in t[3:3 + ntmp]]
        Cd = Cd * 1.0E-6
        Ce = np.zeros_like(Cd)
        for t_ in range(ntmp):
            Ce[t_] = Cd[t_] * weight[icol] / weight[jcol] * np.exp(
                -(energy[icol] - energy[jcol]) / (KB * tmp[t_]))
        colpar.append(dict(name=name, orth_or_para_H2=orth_or_para, ncol=ncol, ntmp=ntmp,
                           icol=icol, jcol=jcol, tmp=tmp, Cd=Cd, Ce=Ce))
    frequency = (energy[irad] - energy[jrad]) / HH
    Bs = A * CC**2 / (2.0 * HH * frequency**3)
    Ba = weight[irad] / weight[jrad] * Bs
    return dict(sym=sym, mass=mass, inverse_mass=float(1.0 / mass), nlev=nlev, energy=energy,
                weight=weight, nrad=nrad, irad=irad, jrad=jrad, A=A, Bs=Bs, Ba=Ba,
                frequency=frequency, colpar=colpar)


def save_linedata(path, ld):
    """Flatten a parsed linedata dict into an .npz fixture."""
    flat = {k: v for k, v in ld.items() if k != "colpar"}
    flat["ncolpar"] = len(ld["colpar"])
    for c, cp in enumerate(ld["colpar"]):
        for k, v in cp.items():
            flat[f"colpar{c}_{k}"] = v
    np.savez_compressed(path, **flat)


def load_linedata(path):
    z = np.load(path, allow_pickle=False)
    ld = {}
    ncolpar = int(z["ncolpar"])
    for k in z.files:
        if k.startswith("colpar") or k == "ncolpar":
            continue
        v = z[k]
        ld[k] = v.item() if v.ndim == 0 else v
    for k in ("nlev", "nrad"):
        ld[k] = int(ld[k])
    ld["sym"] = str(ld["sym"])
    ld["colpar"] = []
    for c in range(ncolpar):
        cp = {}
        pre = f"colpar{c}_"
        for k in z.files:
            if k.startswith(pre):
                v = z[k]
                cp[k[len(pre):]] = v.item() if v.ndim == 0 else v
        cp["name"] = str(cp["name"])
        cp["orth_or_para_H2"] = str(cp["orth_or_para_H2"])
        cp["ncol"] = int(cp["ncol"])
        cp["ntmp"] = int(cp["ntmp"])
        ld["colpar"].append(cp)
    return ld
