#!/usr/bin/env python
"""bench.py -- point*ray*freq Feautrier order-2 solves per second on N B200s (BASELINE.json metric).

Workload (config.workload): BASELINE.json configs[2], the largest configuration that is practical on one GPU inside a
few minutes: synthetic 3D spherically symmetric wind point cloud, 100k points (+ boundary shell), 12*nside^2 HEALPix
rays with nside = 4 (96 ray pairs), CO (41 levels, 40 lines), 11 quadrature points -> 440 frequency bins, LTE level
populations. One "step" = one full `compute_radiation_field_feautrier_order_2` pass (reference src/model/model.cpp:353-390):
ray trace of every (origin, ray) pair, eta/chi assembly, 4.4e9 tridiagonal solves, J and Lambda accumulation; on N GPUs the
origins are partitioned and the J/Lambda slices all-gathered over NCCL (strong scaling).

  value : device-resident inputs, CUDA-event time on the library's stream, max over ranks (no cached ray paths).
  e2e   : the same pass through mgb_solve_feautrier_order_2 with HOST buffers (pinned): H2D of the per-iteration line
          state, run, all-gather, D2H of J and Lambda inside the timed region.
  --impl reference : the unmodified reference (oracle/_ref, compiled from /root/reference by oracle/build_ref.sh) on the
          host cores, same point cloud, a bounded ray-pair subset per step (solves/s is per (o,rr,f) so it compares directly).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--npoints", type=int, default=100000)
    ap.add_argument("--nside", type=int, default=4)
    ap.add_argument("--nquads", type=int, default=11)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-image", action="store_true", help="skip the imaging leg")
    ap.add_argument("--image-pixels", type=int, default=512)
    ap.add_argument("--image-freqs", type=int, default=100)
    ap.add_argument("--no-levelpop", action="store_true", help="skip the level-population iteration leg (s_per_levelpop_iter)")
    ap.add_argument("--ref-pairs", type=int, default=4, help="reference arm: ray pairs spread over the sphere, one per step in turn")
    ap.add_argument("--e2e-steps", type=int, default=3, help="steps of the end-to-end (host buffer) pass")
    ap.add_argument("--levelpop-iters", type=int, default=2)
    ap.add_argument("--no-config3", action="store_true", help="skip the configs[3] leg (1M points, CO + SiO-like, sparse level populations)")
    ap.add_argument("--c3-points", type=int, default=1000000)
    ap.add_argument("--c3-pairs", type=int, default=4, help="ray pairs of the nside-8 set (384) run per radiation-field call of the configs[3] leg")
    ap.add_argument("--no-config4", action="store_true", help="skip the configs[4] leg (10M-point lattice image)")
    ap.add_argument("--c4-points", type=int, default=10000000)
    ap.add_argument("--build-spec", default=None, help=argparse.SUPPRESS)  # internal: background model builder (c3 | c4)
    ap.add_argument("--build-out", default=None, help=argparse.SUPPRESS)
    return ap.parse_args()


# ---------------- background model builders (host-side numpy / Qhull work, overlapped with the GPU legs) ----------------
def spread_pairs(full_h, n):
    """n ray-pair indices spread over the HEALPix ring order (pole to equator), not the first n (all near one pole)."""
    n = max(1, min(n, full_h))
    return sorted({int(round(i * (full_h - 1) / max(1, n - 1))) if n > 1 else 0 for i in range(n)})


def restrict_pairs(spec, pairs):
    full_h = len(spec["direction"]) // 2
    d, w = spec["direction"], spec["weight"]
    idx = list(pairs) + [full_h + p for p in pairs]
    spec["direction"] = np.ascontiguousarray(d[idx])
    spec["weight"] = np.ascontiguousarray(w[idx])
    return full_h


def build_spec_main(args):
    import pickle
    from magritte_b200 import modelbuild as mb
    if args.build_spec == "c3":
        spec = mb.spec_wind_3d(npoints=args.c3_points, nside_rays=8, nquads=args.nquads, spiral=True, species=("co", "sio"), seed=1234)
    else:
        spec = mb.spec_lattice_sphere(npoints=args.c4_points)
    with open(args.build_out + ".tmp", "wb") as f:
        pickle.dump(spec, f, protocol=4)
    os.replace(args.build_out + ".tmp", args.build_out)


class SpecBuilder:
    """Local rank 0 builds the big synthetic point clouds in background processes while the GPU runs the main leg; every rank of the
    node loads the pickle."""

    def __init__(self, args, local_rank):
        self.paths, self.procs = {}, {}
        tag = os.environ.get("MASTER_PORT", "0") + "_" + str(os.getppid() if int(os.environ.get("WORLD_SIZE", "1")) > 1 else os.getpid())
        for name, skip in (("c3", args.no_config3), ("c4", args.no_config4)):
            if skip:
                continue
            self.paths[name] = f"/tmp/mgb_bench_{name}_{tag}.pkl"
            if local_rank == 0:
                cmd = [sys.executable, os.path.abspath(__file__), "--build-spec", name, "--build-out", self.paths[name],
                       "--c3-points", str(args.c3_points), "--c4-points", str(args.c4_points), "--nquads", str(args.nquads)]
                self.procs[name] = subprocess.Popen(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)

    def load(self, name, timeout=900.0):
        import pickle
        t0 = time.perf_counter()
        p = self.procs.get(name)
        while not os.path.exists(self.paths[name]):
            if p is not None and p.poll() is not None and not os.path.exists(self.paths[name]):
                raise RuntimeError(f"builder {name} failed: {p.stderr.read().decode()[-400:]}")
            if time.perf_counter() - t0 > timeout:
                raise RuntimeError(f"builder {name} timed out")
            time.sleep(0.5)
        with open(self.paths[name], "rb") as f:
            return pickle.load(f), time.perf_counter() - t0

    def cleanup(self):
        for p in self.procs.values():
            if p.poll() is None:
                p.kill()
        for path in self.paths.values():
            for q in (path, path + ".tmp"):
                try:
                    os.remove(q)
                except OSError:
                    pass


def workload_spec(args):
    from magritte_b200 import modelbuild as mb
    return mb.spec_wind_3d(npoints=args.npoints, nside_rays=args.nside, nquads=args.nquads, seed=1234)


def config_of(args, flat, n_gpus):
    return {"workload": f"configs[2]: synthetic 3D spherically symmetric wind point cloud, {args.npoints} points + "
                        f"{flat['npoints'] - args.npoints} boundary-shell points, HEALPix nside={args.nside} "
                        f"({flat['hnrays']} ray pairs), CO {flat['nlines']} lines x {flat['nquads']} quadrature points = "
                        f"{flat['nfreqs']} frequency bins, LTE populations, CloseLines approximation, n_off_diag=0",
            "npoints": int(flat["npoints"]), "hnrays": int(flat["hnrays"]), "nfreqs": int(flat["nfreqs"]),
            "solves_per_step": int(flat["npoints"]) * int(flat["hnrays"]) * int(flat["nfreqs"]),
            "partition": f"origins/{n_gpus}", "l2": "inputs larger than L2 (line tables + nu + paths >> 126 MB); "
                                                    "ray paths re-traced every step (no cached paths)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(flat, st_steps, st_nbs):
    """every datum counted once per (o, rr) work item, no reuse across items."""
    N, H, F, L = int(flat["npoints"]), int(flat["hnrays"]), int(flat["nfreqs"]), int(flat["nlines"])
    return 28 * st_nbs + 28 * st_steps + 32 * L * st_steps + 80 * N * H + 16 * N * F + 12 * N * L


def algorithmic_flops(flat, n_entries, n_solves):
    """54 fp64 operations per ray entry per frequency + 110 per solve (default path)."""
    return 54 * int(flat["nquads"]) * n_entries + 110 * n_solves


def cpu_baseline_port(flat, seconds, approx, J_gpu=None):
    """The oracle (CPU restatement of the reference's algorithm, OpenMP over origins) on a bounded origin sample. With J_gpu (the
    full-size J of the CUDA path) the sample doubles as a full-size parity check: the oracle's rows, in the reference's elimination
    order, against the same rows of the GPU result."""
    from oracle import oracle as orc
    cores = os.cpu_count() or 1
    o = orc.Oracle(flat, approx=approx, store_u=False, threads=cores)
    N, H, F = int(flat["npoints"]), int(flat["hnrays"]), int(flat["nfreqs"])
    rng = np.random.default_rng(0)
    start = int(rng.integers(0, max(1, N - 4096)))
    hl = 4096  # scratch half-length (rays of this workload are < 600 entries)
    n0 = min(N, max(cores * 4, 64))
    t0 = time.perf_counter()
    o.solve(o_range=(start, start + n0), rr_range=(0, 1), half_length=hl)
    pilot = time.perf_counter() - t0
    per_o_rr = pilot / n0
    # sample: `no` consecutive origins x all ray pairs, sized to ~`seconds`
    no = int(max(cores * 2, min(N - start, seconds / max(per_o_rr * H, 1e-9))))
    t0 = time.perf_counter()
    o.solve(o_range=(start, start + no), rr_range=(0, H), half_length=hl)
    dt = time.perf_counter() - t0
    solves = no * H * F
    res = {"value": solves / dt, "unit": "solves/s", "cores": cores, "kind": "port",
           "sample": f"{no} consecutive origins (from {start}) x all {H} ray pairs x {F} freqs = {solves} solves in {dt:.2f} s, "
                     f"oracle/feautrier_oracle.c with OpenMP"}
    if J_gpu is not None:
        a, b = np.asarray(J_gpu[start:start + no]), o.J[start:start + no]
        res["parity_vs_gpu"] = {"origins": no, "max_rel_err_J": float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))),
                                "tolerance": 1e-10, "note": "full-size workload: oracle rows (reference elimination order) vs the CUDA result"}
    return res


def run_reference(args):
    """--impl reference: the compiled, unmodified reference on the host cores (rank 0 only). Each step is ONE ray pair of the configs[2]
    cloud (a bounded sample: the work of an (o, rr, f) solve does not depend on how many pairs a call runs), and the steps take
    --ref-pairs different pairs spread over the HEALPix sphere in turn (one model per pair, built before the timed region). Also
    timed, outside the K steps: the reference's own level-population iteration (its timer_1 + timer_2, model.cpp:706-724) on configs[1]
    in full and on the configs[2] cloud with one ray pair."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import refcore
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import util
    from magritte_b200 import modelbuild as mb
    spec0 = workload_spec(args)
    full_h = len(spec0["direction"]) // 2
    pairs = spread_pairs(full_h, args.ref_pairs)
    cores = os.cpu_count() or 1
    N = len(spec0["position"])
    nfreqs = sum(int(ld["nrad"]) for ld in spec0["linedata"]) * args.nquads
    cfg = {"workload": f"configs[2] point cloud ({args.npoints} points + shell), CO {nfreqs} frequency bins; each step of the reference runs "
                       f"ONE of the {full_h} ray pairs (bounded sample; work per (o,rr,f) solve is independent of the number of pairs), "
                       f"taking pairs {pairs} (spread over the HEALPix ring order) in turn",
           "npoints": N, "hnrays": 1, "nfreqs": nfreqs, "solves_per_step": N * nfreqs, "ray_pairs_cycled": pairs}
    levelpop = None
    if refcore.available():
        core, _ = refcore.load()
        core.pcmt_set_n_threads_avail(cores)
        models = []
        for i, pr in enumerate(pairs):
            spec = dict(spec0)
            restrict_pairs(spec, [pr])
            with util.quiet():
                m = mb.build(core, core.IoPython("mem", f"bench_ref{i}"), f"bench_ref{i}", spec)
                mb.prepare(m)
            models.append(m)
        times = []
        for it in range(args.warmup + args.steps):
            m = models[it % len(models)]
            t0 = time.perf_counter()
            with util.quiet():
                m.compute_radiation_field_feautrier_order_2()
            if it >= args.warmup:
                times.append(time.perf_counter() - t0)
        kind, how = "reference", "oracle/_ref/magritte/core.so (unmodified Magritte 0.9.3, g++ -O3 -fopenmp), compute_radiation_field_feautrier_order_2"
        # ---- the reference's seconds per level-population iteration ----
        levelpop = {}
        try:
            c2 = mb.spec_vanzadelhoff_1a_1d()
            with util.quiet():
                m2 = mb.build(core, core.IoPython("mem", "bench_ref_c2"), "bench_ref_c2", c2)
                mb.prepare(m2)
                m2.parameters.use_adaptive_Ng_acceleration = False
                t0 = time.perf_counter()
                n2 = m2.compute_level_populations(False, 5)
                dt2 = time.perf_counter() - t0
            levelpop["configs[1] van Zadelhoff 1a (250 points, 100 ray pairs, full)"] = {"value": dt2 / max(n2, 1), "unit": "s/iteration", "iterations": int(n2)}
        except Exception as e:  # noqa: BLE001
            levelpop["configs[1]"] = {"error": repr(e)[:200]}
        if not args.no_levelpop:
            try:
                m0 = models[0]
                m0.parameters.use_adaptive_Ng_acceleration = False
                t0 = time.perf_counter()
                with util.quiet():
                    n0 = m0.compute_level_populations(False, 1)
                dt0 = time.perf_counter() - t0
                rf = float(np.mean(times))  # seconds of one radiation-field call with one ray pair (the timed steps above)
                levelpop[f"configs[2] cloud, ONE of the {full_h} ray pairs"] = {
                    "value": dt0 / max(n0, 1), "unit": "s/iteration", "iterations": int(n0),
                    "radiation_field_one_pair_s": rf, "statistical_equilibrium_s": dt0 / max(n0, 1) - rf,
                    "extrapolated_full_iteration_s": rf * full_h + (dt0 / max(n0, 1) - rf),
                    "note": "the reference's global SparseLU statistical equilibrium is serial (lineProducingSpecies.tpp:377-581); the full iteration is "
                            "extrapolated as radiation_field_one_pair x hnrays + statistical equilibrium and labelled as such (SURVEY.md 8d)"}
            except Exception as e:  # noqa: BLE001
                levelpop["configs[2]"] = {"error": repr(e)[:200]}
    else:
        from magritte_b200 import synth
        from oracle import oracle as orc
        flats = []
        for pr in pairs:
            spec = dict(spec0)
            restrict_pairs(spec, [pr])
            flats.append(synth.flat_from_spec(spec))
        oracles = [orc.Oracle(f, store_u=False, threads=cores) for f in flats]
        times = []
        for it in range(args.warmup + args.steps):
            o = oracles[it % len(oracles)]
            t0 = time.perf_counter()
            o.solve(half_length=4096)
            if it >= args.warmup:
                times.append(time.perf_counter() - t0)
        kind, how = "port", "oracle/feautrier_oracle.c (CPU restatement, OpenMP); oracle/_ref not present"
    ms = 1e3 * float(np.mean(times))
    solves = N * nfreqs
    val = solves / (ms * 1e-3)
    out = {"impl": "reference", "metric": "point*ray*freq Feautrier order-2 solves/s", "value": val, "unit": "solves/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
           "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": cfg, "cpu_baseline": {"value": val, "unit": "solves/s", "cores": cores, "kind": kind,
                                           "sample": f"1 ray pair x {N} origins x {nfreqs} freqs per step, pairs {pairs} in turn; {how}"},
           "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "s_per_levelpop_iter": levelpop}
    emit(out)


def oracle_rows(flat_geom, m, start, count, pairs_h, sparse_weights=None):
    """Parity sample of a host-Model leg: the oracle (reference elimination order) on `count` origins with the MODEL'S current line
    state, against the same rows of the model's J (or, for the memory-sparse solver, of lspec.J = sum_z w_z J)."""
    from oracle import oracle as orc
    N = int(flat_geom["npoints"])
    f2 = dict(flat_geom)
    L = int(f2["nlines"])
    nq = int(f2["nquads"])
    f2["nfreqs"] = L * nq
    nu = np.zeros((N, L * nq))
    nu[start:start + count] = m.get_nu_rows(start, start + count)
    f2["nu"] = nu
    f2["corresponding_line"] = (np.arange(L * nq, dtype=np.uint32) // nq).astype(np.uint32)
    f2["emissivity"], f2["opacity"], f2["inverse_width"] = m.get_emissivity(), m.get_opacity(), m.get_inverse_width()
    f2["population_tot"] = np.ascontiguousarray(np.stack([m.get_population_tot(l) for l in range(int(f2["nlspecs"]))]))
    o = orc.Oracle(f2, store_u=False, threads=os.cpu_count() or 1)
    o.solve(o_range=(start, start + count), rr_range=(0, pairs_h), half_length=8192)
    want = o.J[start:start + count]
    if sparse_weights is None:
        got = m.get_J_rows(start, start + count)
    else:
        want = (want.reshape(count, L, nq) * sparse_weights[None]).sum(axis=2)
        got = np.concatenate([m.get_J_species(l)[start:start + count] for l in range(int(f2["nlspecs"]))], axis=1)
    return float(np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)))


def host_model(spec, flat_geom, device, world, rank, dist, **parameters):
    from magritte_b200 import model as hostmodel
    m = hostmodel.from_spec(spec, flat_geom, device, **parameters)
    if world > 1:
        uid = [m.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        m.set_communicator(world, rank, uid[0])
    return m


def levelpop_leg(args, spec, flat_geom, device, world, rank, dist, sparse, iters, label, pairs_h, parity_origins=48):
    """BASELINE.json's second metric: seconds per level-population iteration = radiation field + compute_Jeff + statistical
    equilibrium (the reference's timer_1 + timer_2, model.cpp:706-724), through the host Model (magritte_b200.core) on `world` GPUs:
    origins partitioned over the ranks, J / Lambda all-gathered over NCCL, level-population half on the device (mgb_lp_*)."""
    t_setup = time.perf_counter()
    m = host_model(spec, flat_geom, device, world, rank, dist, use_adaptive_Ng_acceleration=False)
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    t_setup = time.perf_counter() - t_setup
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    n = m.compute_level_populations_sparse(False, iters) if sparse else m.compute_level_populations(False, iters)
    wall = time.perf_counter() - t0
    first = dict(m.last_stats)
    # one more radiation field at the iterated populations (cached paths): the per-call time an ALI loop sees from iteration 2 on, and
    # the state the parity sample is taken at
    t1 = time.perf_counter()
    if sparse:
        m.compute_radiation_field_feautrier_order_2_sparse()
    else:
        m.compute_radiation_field_feautrier_order_2()
    t_rf = time.perf_counter() - t1
    N = int(flat_geom["npoints"])
    res = {"workload": label, "value": wall / max(n, 1), "unit": "s/iteration", "iterations": int(n), "n_gpus": world,
           "radiation_field_and_Jeff_s": m.seconds_radiation_field / max(n, 1),
           "statistical_equilibrium_s": m.seconds_statistical_equilibrium / max(n, 1),
           "radiation_field_host_buffers_s": t_rf, "host_setup_s": t_setup, "origins_of_rank0": [int(m.o_begin), int(m.o_end)],
           "solves_per_iteration": N * pairs_h * int(flat_geom["nlines"]) * int(flat_geom["nquads"]),
           "error_max": [float(x) for x in m.error_max],
           "note": "level-population half on the device (Jeff, statistical equilibrium in double-double, line-state refresh): no per-point table "
                   "crosses PCIe inside the loop; the first iteration includes the ray trace; N > 1: cost-balanced origin partition + NCCL all-gather"}
    if rank == 0:
        start = N // 3
        nq = int(flat_geom["nquads"])
        w = np.repeat(np.asarray(flat_geom["quad_weights"]), np.asarray(flat_geom["nrad"]), axis=0) if sparse else None
        err = oracle_rows(flat_geom, m, start, parity_origins, pairs_h, sparse_weights=w)
        res["parity"] = {"origins": parity_origins, "max_rel_err": err, "tolerance": 1e-10,
                         "what": ("lspec.J (line-integrated J)" if sparse else "J") + " rows of the radiation field at the iterated populations vs the oracle "
                                 "(reference elimination order, same line state)"}
        if err > 1e-10:
            raise SystemExit(f"bench.py: parity sample of the {label} leg failed: {res['parity']}")
    del m
    return res


def image_leg(args, spec, flat_geom, device, world, rank, dist, label):
    """BASELINE configs[4]: synthetic observation of the 10M-point cloud, 512 x 512 pixels x 100 frequency bins around CO J=2-1, pixels
    partitioned over the ranks, through Model.compute_image_new (host buffers: D2H of the image inside the timed region)."""
    t_setup = time.perf_counter()
    m = host_model(spec, flat_geom, device, world, rank, dist)
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    nu0 = float(flat_geom["line"][0])
    dv = 20.0e3 / 2.99792458e8
    m.compute_spectral_discretisation(nu0 * (1 - dv), nu0 * (1 + dv), args.image_freqs)
    nx = ny = args.image_pixels
    m.compute_image_new(0.3, 0.5, 0.81, nx, ny)  # warm-up: uploads the model, sizes the device buffers of this image shape
    t_setup = time.perf_counter() - t_setup
    times, kern = [], []
    for d in [(1.0, 0.0, 0.0), (0.3, 0.5, 0.81)]:
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        m.compute_image_new(d[0], d[1], d[2], nx, ny)
        times.append(time.perf_counter() - t0)
        kern.append(dict(m.last_image_stats))
    res = {"workload": label, "ms_per_image": 1e3 * float(np.mean(times)), "ms_each": [1e3 * t for t in times], "pixels": [nx, ny],
           "nfreqs": args.image_freqs, "n_gpus": world, "kernel_ms_rank0": [{"trace": k["ms_trace"], "solve": k["ms_solve"], "total": k["ms_total"]} for k in kern],
           "host_setup_s": t_setup, "npoints": int(flat_geom["npoints"])}
    if rank == 0:
        from oracle import oracle as orc
        nu = np.asarray(m.images[-1].freqs)
        f2 = dict(flat_geom)
        N = int(f2["npoints"])
        f2["nfreqs"] = len(nu)
        f2["nu"] = np.ascontiguousarray(np.broadcast_to(nu, (N, len(nu))))
        f2["corresponding_line"] = np.zeros(len(nu), dtype=np.uint32)
        f2["emissivity"], f2["opacity"], f2["inverse_width"] = m.get_emissivity(), m.get_opacity(), m.get_inverse_width()
        f2["population_tot"] = np.ascontiguousarray(np.stack([m.get_population_tot(l) for l in range(int(f2["nlspecs"]))]))
        want = orc.Oracle(f2, store_u=False, threads=os.cpu_count() or 1).image((0.3, 0.5, 0.81), 32, 32, 3)
    # parity sample: a 32 x 32 image of the same model and direction against the oracle (rank 0's share of the pixels when partitioned)
    m.compute_image_new(0.3, 0.5, 0.81, 32, 32)
    if rank != 0:
        return res
    im = m.images[-1]
    I = np.asarray(im.I)
    p0, p1 = int(im.pix_begin), int(im.pix_end)
    a, b = I[p0:p1], want["I"][p0:p1]
    err = float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.0
    res["parity"] = {"what": "32 x 32 image of the same model and direction vs the oracle" + (" (rank 0's pixel share)" if world > 1 else ""),
                     "bit_identical": bool(np.array_equal(a, b)), "max_rel_err": err, "tolerance": 1e-10,
                     "ImX_equal": bool(np.array_equal(np.asarray(im.ImX), want["ImX"]))}
    if err > 1e-10:
        raise SystemExit(f"bench.py: parity sample of the image leg failed: {res['parity']}")
    return res


def run_b200(args):
    import torch
    from magritte_b200 import cabi, synth
    from magritte_b200.flatten import approx_of
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    builder = SpecBuilder(args, local)  # configs[3] / configs[4] point clouds: built in the background while the GPU runs the main leg
    try:
        out = run_b200_legs(args, torch, dist, cabi, synth, approx_of, rank, world, local, builder)
    finally:
        builder.cleanup()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        emit(out)


def run_b200_legs(args, torch, dist, cabi, synth, approx_of, rank, world, local, builder):
    spec_main = workload_spec(args)
    flat = synth.flat_from_spec(spec_main)
    approx = approx_of(flat["params"])
    ctx = cabi.Context(local)
    ctx.set_model(flat)
    if world > 1:
        uid = [ctx.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(world, rank, uid[0])
    o_range = ctx.partition(world, rank, balanced=True)
    N, H, F, L = int(flat["npoints"]), int(flat["hnrays"]), int(flat["nfreqs"]), int(flat["nlines"])
    solves_total = N * H * F

    def sync():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- device-resident pass (value) ----------------
    def step_device():
        ctx.invalidate_paths()  # full work every step: count pass + trace + solve + reduce (+ all-gather)
        st = ctx.run(approx, o_range)
        if world > 1:
            ctx.allgather()
        return st

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local)
    sync()
    if rank == 0:
        sampler.start()
    ctx.timer_start()
    stats = [step_device() for _ in range(args.steps)]
    ms_dev = ctx.timer_stop()
    sync()
    ms_dev = max_over_ranks(ms_dev)
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms_dev / args.steps
    value = solves_total / (ms_per_step * 1e-3)
    st = stats[-1]
    launches = sum(s.n_launches for s in stats)

    # cached-path variant (what an ALI loop sees after its first iteration), informational
    ctx.run(approx, o_range)
    sync()
    ctx.timer_start()
    ctx.run(approx, o_range)
    if world > 1:
        ctx.allgather()
    ms_cached = max_over_ranks(ctx.timer_stop())

    # ---------------- end-to-end pass through the host-buffer call (e2e) ----------------
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()
    h_e, h_o, h_w, h_t = (pin(flat[k]) for k in ("emissivity", "opacity", "inverse_width", "population_tot"))
    no = o_range[1] - o_range[0]
    h_J = pin(np.zeros((N if world > 1 else no, F)))
    h_lv = pin(np.zeros((N if world > 1 else no, L)))
    h_lc = pin(np.zeros((N if world > 1 else no, L), dtype=np.uint32))

    def step_e2e():
        ctx.invalidate_paths()
        if world == 1:
            ctx.solve(h_e, h_o, h_w, h_t, approx=approx, o_range=o_range, J=h_J, lam=(h_lv, h_lc))
        else:
            ctx.set_line_state(h_e, h_o, h_w, h_t)
            ctx.run(approx, o_range)
            ctx.allgather()
            import ctypes as C
            ctx._check(ctx.lib.mgb_get_J(ctx._ctx, C.c_uint32(0), C.c_uint32(N), C.c_void_p(h_J.ctypes.data)), "mgb_get_J")
            ctx._check(ctx.lib.mgb_get_lambda(ctx._ctx, C.c_uint32(0), C.c_uint32(N), C.c_void_p(h_lv.ctypes.data),
                                              C.c_void_p(h_lc.ctypes.data)), "mgb_get_lambda")
        return float(h_J[0, 0])

    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    step_e2e()
    sync()
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    ms_e2e_dev = ctx.timer_stop()
    ms_e2e_wall = (time.perf_counter() - t0) * 1e3
    sync()
    ms_e2e = max_over_ranks(max(ms_e2e_dev, ms_e2e_wall)) / e2e_steps
    h2d = int(h_e.nbytes + h_o.nbytes + h_w.nbytes + h_t.nbytes)
    d2h = int(h_J.nbytes + h_lv.nbytes + h_lc.nbytes)

    # ---------------- roofline of the dominant kernel (k_solve): max(t_HBM, t_FP64), SURVEY.md 8(d) ----------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    ms_solve = float(np.mean([s.ms_solve for s in stats]))
    n_solve_launches = int(np.mean([s.n_solve_launches for s in stats]))
    share = no / N if world > 1 else 1.0
    B = algorithmic_bytes(flat, st.n_steps, st.n_nbs_scanned) * share  # this rank's launches
    Fl = algorithmic_flops(flat, st.n_entries, st.n_solves)
    fp64_peak = ctx.fp64_peak_tflops()
    t_hbm = B / (hbm_peak * 1e9)
    t_fp64 = Fl / (fp64_peak * 1e12) if fp64_peak else 0.0
    pinfo = ctx.pipeline_info()
    pipeline = pinfo["kernel"] == 4
    traffic, traffic_note = None, "not captured for this configuration"
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        key = "pipeline" if pipeline else "k_solve"
        t = tj.get(key, tj if not pipeline else {})
        if world == 1 and args.npoints == 100000 and args.nside == 4 and args.nquads == 11 and t.get("dram_bytes_per_step"):
            traffic, traffic_note = t["dram_bytes_per_step"] / max(1, pinfo["n_ranges"] if pipeline else n_solve_launches), t.get("note", "")
    except Exception:
        pass
    if pipeline:
        n_solve_launches = max(1, pinfo["n_ranges"])  # one k_fill / k_large / k_sweep launch per record range
    hbm_line = {"bound": "hbm", "achieved": B / (ms_solve * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": B / (ms_solve * 1e-3) / 1e9 / hbm_peak,
                "peak_source": peak_src, "roofline_time_ms": 1e3 * t_hbm}
    fp64_line = {"bound": "fp64", "achieved": Fl / (ms_solve * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                 "frac": Fl / (ms_solve * 1e-3) / 1e12 / fp64_peak if fp64_peak else None,
                 "peak_source": "mgb_measure_fp64_peak: a DFMA loop measured on this GPU in this run, 2 flop per DFMA (MEASURED_PEAKS.json holds no FP64 "
                                "figure; nominal B200 FP64 is 37-40 TFLOP/s)", "roofline_time_ms": 1e3 * t_fp64}
    roofline = dict(fp64_line if t_fp64 >= t_hbm else hbm_line)
    kernel_name = ("solve pipeline: k_ni + k_fill (ray-entry records) -> k_large (large-shift line integrals) -> k_sweep<CloseLines> "
                   "(eta/chi + two-sided elimination), one launch of each per record range") if pipeline else "k_solve<CloseLines,false>"
    roofline.update({"kernel": kernel_name, "traffic": traffic, "traffic_note": traffic_note,
                     "algorithmic_bytes_per_launch": B / max(1, n_solve_launches), "algorithmic_flops_per_launch": Fl / max(1, n_solve_launches),
                     "avg_launch_ms": ms_solve / max(1, n_solve_launches), "launches_per_step": n_solve_launches,
                     "flop_model": "54*nquads*SigmaN + 110*solves (SURVEY.md 8d)",
                     "byte_model": "28*SigmaD + 28*SigmaS + 32*nlines*SigmaS + 80*N*H + 16*N*F + 12*N*L (SURVEY.md 8d)",
                     "other_bound": hbm_line if t_fp64 >= t_hbm else fp64_line,
                     "note": "roofline time = max(t_HBM, t_FP64) (SURVEY.md 8d): the FP64 pipe bounds this kernel; frac = roofline time / measured time. "
                             "N > 1: rank 0's launches"})
    if pipeline:
        roofline["pipeline_ms"] = {"records": pinfo["ms_records"], "line_integrals": pinfo["ms_line_integrals"], "sweep": pinfo["ms_sweep"],
                                   "record_ranges": pinfo["n_ranges"], "note": "device time of the three stages in the last timed step (CUDA events)"}
    out = None
    if rank == 0:
        out = {"metric": "point*ray*freq Feautrier order-2 solves/s", "value": value, "unit": "solves/s", "n_gpus": world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_of(args, flat, world),
               "clocks": clocks, "e2e": {"value": solves_total / (ms_e2e * 1e-3), "unit": "solves/s", "h2d_bytes_per_step": h2d,
                                         "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e, "steps": e2e_steps},
               "gpu_launches": int(launches), "roofline": roofline, "roofline_fp64": fp64_line,
               "breakdown_ms": {"trace_count": float(np.mean([s.ms_trace_count for s in stats])),
                                "trace_fill": float(np.mean([s.ms_trace_fill for s in stats])), "solve": ms_solve,
                                "reduce": float(np.mean([s.ms_reduce for s in stats])),
                                "run_total": float(np.mean([s.ms_total for s in stats]))},
               "work": {"sigma_S_steps": int(st.n_steps), "sigma_D_neighbours": int(st.n_nbs_scanned), "sigma_N_entries": int(st.n_entries),
                        "large_shift_evaluations": int(st.n_large_shift), "max_ray_length": int(st.max_ray_length),
                        "origin_chunks": int(st.n_chunks)},
               "cached_paths": {"value": solves_total / (ms_cached * 1e-3), "unit": "solves/s", "ms_per_step": ms_cached,
                                "note": "ray paths kept in HBM across calls (static geometry), as inside an ALI loop"},
               "s_per_levelpop_iter": None}
        if not args.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline_port(flat, args.cpu_seconds, approx, J_gpu=h_J)
            if out["cpu_baseline"]["parity_vs_gpu"]["max_rel_err_J"] > 1e-10:
                raise SystemExit(f"bench.py: full-size parity check failed: {out['cpu_baseline']['parity_vs_gpu']}")
        elif world == 1:
            out["cpu_baseline"] = None
    # ---------------- imaging leg on the bench model: pixels partitioned over the ranks ----------------
    if not args.no_image:
        nx = ny = args.image_pixels
        nu0 = float(flat["line"][1])  # CO J=2-1
        dv = 20.0e3 / 2.99792458e8
        nu_img = np.linspace(nu0 * (1 - dv), nu0 * (1 + dv), args.image_freqs)
        npix = nx * ny
        p0, p1 = npix * rank // world, npix * (rank + 1) // world
        ctx.image(2, 3, (0.3, 0.5, 0.81), nx, ny, nu_img, pix_range=(p0, p1))  # warm-up
        sync()
        ctx.timer_start()
        t0 = time.perf_counter()
        img_each = []
        for d in [(1.0, 0.0, 0.0), (0.3, 0.5, 0.81)]:
            t1 = time.perf_counter()
            _, _, I_img, ist = ctx.image(2, 3, d, nx, ny, nu_img, pix_range=(p0, p1))
            img_each.append({"wall_ms": (time.perf_counter() - t1) * 1e3, "call_ms": ist.ms_total, "trace": ist.ms_trace, "solve": ist.ms_solve})
        ms_img = max(ctx.timer_stop(), (time.perf_counter() - t0) * 1e3)
        ms_img = max_over_ranks(ms_img) / 2
        if out is not None:
            out["image"] = {"metric": "pixel*freq image solves/s (Solver::image_feautrier_order_2_new_imager)", "value": npix * args.image_freqs / (ms_img * 1e-3),
                            "unit": "solves/s", "ms_per_image": ms_img, "pixels": [nx, ny], "nfreqs": args.image_freqs,
                            "kernel_ms": {"trace": ist.ms_trace, "solve": ist.ms_solve}, "each": img_each, "entries_per_pixel": ist.n_entries / max(1, ist.n_pixels),
                            "note": "end to end through mgb_image_feautrier_order_2 with host buffers (D2H of the image inside the timed region); "
                                    f"the {flat['npoints']}-point model of this bench (configs[4]'s 10M-point cloud: image_10m); pixels/{world} per rank"}
    ctx.close()
    del h_J, h_lv, h_lc, h_e, h_o, h_w, h_t
    torch.cuda.empty_cache()

    def leg(fn, *a, **kw):
        """A secondary leg must not take the headline down with it: its failure is reported in its own key."""
        try:
            return fn(*a, **kw)
        except SystemExit:
            raise
        except Exception as e:  # noqa: BLE001
            import traceback
            return {"error": repr(e)[:300], "trace": traceback.format_exc()[-600:]}

    # ---------------- second metric: seconds per level-population iteration, configs[2], N GPUs ----------------
    if not args.no_levelpop:
        geom = synth.flat_from_spec(spec_main, with_state=False)
        r = leg(levelpop_leg, args, spec_main, geom, local, world, rank, dist, False, args.levelpop_iters,
                f"configs[2]: {N} points, {H} ray pairs, {F} frequency bins, compute_level_populations(False, {args.levelpop_iters})", H)
        if out is not None:
            out["s_per_levelpop_iter"] = r
    # ---------------- configs[3]: 1M points, nside 8 (ray-pair subset), CO + SiO-like, memory-sparse level populations ----------------
    if not args.no_config3:
        def c3():
            spec, waited = builder.load("c3")
            full_h = restrict_pairs(spec, spread_pairs(len(spec["direction"]) // 2, args.c3_pairs))
            geom = synth.flat_from_spec(spec, with_state=False)
            Hc = int(geom["hnrays"])
            label = (f"configs[3]: {int(geom['npoints'])} points (spiral wind), CO + SiO-like = {int(geom['nlines'])} lines x {args.nquads} = "
                     f"{int(geom['nlines']) * args.nquads} frequency bins, {Hc} of the {full_h} ray pairs of HEALPix nside 8 (spread over the sphere; the "
                     f"time box does not hold 384), compute_level_populations_sparse(False, {args.levelpop_iters})")
            r = levelpop_leg(args, spec, geom, local, world, rank, dist, True, args.levelpop_iters, label, Hc)
            r["builder_wait_s"] = waited
            r["solves_per_s"] = r["solves_per_iteration"] / r["radiation_field_and_Jeff_s"]
            r["extrapolated_s_per_iteration_all_ray_pairs"] = r["radiation_field_and_Jeff_s"] * full_h / Hc + r["statistical_equilibrium_s"]
            return r
        r = leg(c3)
        if out is not None:
            out["config3"] = r
    # ---------------- configs[4]: 512 x 512 x 100 image of the 10M-point cloud, pixels / N per GPU ----------------
    if not args.no_config4:
        def c4():
            spec, waited = builder.load("c4")
            geom = synth.flat_from_spec(spec, with_state=False)
            r = image_leg(args, spec, geom, local, world, rank, dist,
                          f"configs[4]: jittered-lattice sphere, {int(geom['npoints'])} points, CO J=2-1, {args.image_pixels}^2 pixels x {args.image_freqs} bins, "
                          f"Model.compute_image_new, pixels/{world} per GPU")
            if r is not None:
                r["builder_wait_s"] = waited
                r["value"] = args.image_pixels ** 2 * args.image_freqs / (r["ms_per_image"] * 1e-3)
                r["unit"] = "pixel*freq solves/s"
            return r
        r = leg(c4)
        if out is not None:
            out["image_10m"] = r
    return out


_STDOUT_FD = None


def emit(obj):
    line = (json.dumps(obj) + "\n").encode()
    if _STDOUT_FD is None:
        sys.stdout.write(line.decode()); sys.stdout.flush()
    else:
        os.write(_STDOUT_FD, line)


def main():
    # rank 0 prints exactly ONE line on stdout. Libraries write there too (NCCL prints "NCCL version ..." from C, whatever NCCL_DEBUG_FILE
    # says): file descriptor 1 is pointed at stderr for the whole run and the JSON line goes to the saved descriptor (emit()).
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    args = parse()
    if args.build_spec:
        return build_spec_main(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
