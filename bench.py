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
    ap.add_argument("--ref-pairs", type=int, default=1, help="ray pairs per step of the reference arm")
    return ap.parse_args()


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
    """--impl reference: the compiled, unmodified reference on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import refcore
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import util
    from magritte_b200 import modelbuild as mb
    spec = workload_spec(args)
    full_h = len(spec["direction"]) // 2
    R = max(1, min(args.ref_pairs, full_h))
    d = spec["direction"]
    spec["direction"] = np.concatenate([d[:R], d[full_h:full_h + R]], axis=0)
    spec["weight"] = np.concatenate([spec["weight"][:R], spec["weight"][full_h:full_h + R]])
    cores = os.cpu_count() or 1
    N = len(spec["position"])
    nfreqs = sum(int(ld["nrad"]) for ld in spec["linedata"]) * args.nquads
    cfg = {"workload": f"configs[2] point cloud ({args.npoints} points + shell), CO {nfreqs} frequency bins; the reference runs "
                       f"{R} of the {full_h} ray pairs per step (bounded sample; work per (o,rr,f) solve is independent of the "
                       f"number of pairs)", "npoints": N, "hnrays": R, "nfreqs": nfreqs, "solves_per_step": N * R * nfreqs}
    if refcore.available():
        core, _ = refcore.load()
        core.pcmt_set_n_threads_avail(cores)
        with util.quiet():
            m = mb.build(core, core.IoPython("mem", "bench_ref"), "bench_ref", spec)
            mb.prepare(m)
        times = []
        for it in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            with util.quiet():
                m.compute_radiation_field_feautrier_order_2()
            if it >= args.warmup:
                times.append(time.perf_counter() - t0)
        kind, how = "reference", "oracle/_ref/magritte/core.so (unmodified Magritte 0.9.3, g++ -O3 -fopenmp), compute_radiation_field_feautrier_order_2"
    else:
        from magritte_b200 import synth
        from oracle import oracle as orc
        flat = synth.flat_from_spec(spec)
        o = orc.Oracle(flat, store_u=False, threads=cores)
        times = []
        for it in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            o.solve(half_length=4096)
            if it >= args.warmup:
                times.append(time.perf_counter() - t0)
        kind, how = "port", "oracle/feautrier_oracle.c (CPU restatement, OpenMP); oracle/_ref not present"
    ms = 1e3 * float(np.mean(times))
    solves = N * R * nfreqs
    val = solves / (ms * 1e-3)
    emit(({"impl": "reference", "metric": "point*ray*freq Feautrier order-2 solves/s", "value": val, "unit": "solves/s",
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": cfg, "cpu_baseline": {"value": val, "unit": "solves/s", "cores": cores, "kind": kind,
                                                      "sample": f"{R} ray pair(s) x {N} origins x {nfreqs} freqs per step; {how}"},
                      "e2e": {"value": val, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "gpu_launches": 0}))


def levelpop_iteration(args, flat, device):
    """BASELINE.json's second metric: seconds per level-population iteration = radiation field + compute_Jeff + statistical
    equilibrium (the reference's timer_1 + timer_2, model.cpp:706-724), through the host layer (magritte_b200.core.Model)."""
    from magritte_b200 import model as hostmodel
    spec = workload_spec(args)
    m = hostmodel.from_spec(spec, flat, device, use_adaptive_Ng_acceleration=False)
    m.compute_spectral_discretisation()
    m.compute_inverse_line_widths()
    m.compute_LTE_level_populations()
    assert np.array_equal(m.get_emissivity(), flat["emissivity"]) and np.array_equal(m.get_nu(), flat["nu"])
    t0 = time.perf_counter()
    n = m.compute_level_populations(False, 2)
    wall = time.perf_counter() - t0
    return {"value": wall / n, "unit": "s/iteration", "iterations": n, "radiation_field_and_Jeff_s": m.seconds_radiation_field / n,
            "statistical_equilibrium_s": m.seconds_statistical_equilibrium / n,
            "note": "host-buffer path: H2D of the line state and D2H of J/Lambda every iteration; first iteration includes the ray trace; "
                    "statistical equilibrium on the host cores (long double, OpenMP)"}


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

    flat = synth.flat_from_spec(workload_spec(args))
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
    stc = ctx.run(approx, o_range)
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

    step_e2e()
    sync()
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    ms_e2e_dev = ctx.timer_stop()
    ms_e2e_wall = (time.perf_counter() - t0) * 1e3
    sync()
    ms_e2e = max_over_ranks(max(ms_e2e_dev, ms_e2e_wall)) / args.steps
    h2d = int(h_e.nbytes + h_o.nbytes + h_w.nbytes + h_t.nbytes)
    d2h = int(h_J.nbytes + h_lv.nbytes + h_lc.nbytes)

    # ---------------- roofline of the dominant kernel (k_solve) ----------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    ms_solve = float(np.mean([s.ms_solve for s in stats]))
    n_solve_launches = int(np.mean([s.n_solve_launches for s in stats]))
    B = algorithmic_bytes(flat, st.n_steps, st.n_nbs_scanned) * (no / N if world > 1 else 1.0)
    Fl = algorithmic_flops(flat, st.n_entries, st.n_solves)
    fp64_peak = ctx.fp64_peak_tflops()
    achieved = B / (ms_solve * 1e-3) / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("k_solve_dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "k_solve<CloseLines>", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": B / max(1, n_solve_launches), "avg_launch_ms": ms_solve / max(1, n_solve_launches),
                "launches_per_step": n_solve_launches,
                "note": "this kernel is FP64-pipe bound, not HBM bound (SURVEY.md 8d): see roofline_fp64"}
    roofline_fp64 = {"bound": "fp64", "achieved": Fl / (ms_solve * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                     "frac": Fl / (ms_solve * 1e-3) / 1e12 / fp64_peak if fp64_peak else None,
                     "peak_source": "mgb_measure_fp64_peak: DFMA loop measured in this run (2 flop per DFMA)",
                     "algorithmic_flops": Fl, "flop_model": "54*nquads*SigmaN + 110*solves (SURVEY.md 8d)"}
    out = None
    if rank == 0:
        out = {"metric": "point*ray*freq Feautrier order-2 solves/s", "value": value, "unit": "solves/s", "n_gpus": world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_of(args, flat, world),
               "clocks": clocks, "e2e": {"value": solves_total / (ms_e2e * 1e-3), "unit": "solves/s", "h2d_bytes_per_step": h2d,
                                         "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e},
               "gpu_launches": int(launches), "roofline": roofline, "roofline_fp64": roofline_fp64,
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
    # ---------------- imaging leg (BASELINE.json configs[4] shape on this model): pixels partitioned over the ranks ----------------
    if not args.no_image:
        nx = ny = args.image_pixels
        nu0 = float(flat["line"][1])  # CO J=2-1
        dv = 20.0e3 / 2.99792458e8
        nu_img = np.linspace(nu0 * (1 - dv), nu0 * (1 + dv), args.image_freqs)
        npix = nx * ny
        p0, p1 = npix * rank // world, npix * (rank + 1) // world
        ctx.image(2, 3, (0.3, 0.5, 0.81), nx, ny, nu_img, pix_range=(p0, p1))  # warm-up
        sync()
        if dist is not None:
            dist.barrier()
        ctx.timer_start()
        t0 = time.perf_counter()
        for d in [(1.0, 0.0, 0.0), (0.3, 0.5, 0.81)]:
            _, _, I_img, ist = ctx.image(2, 3, d, nx, ny, nu_img, pix_range=(p0, p1))
        ms_img = max(ctx.timer_stop(), (time.perf_counter() - t0) * 1e3)
        ms_img = max_over_ranks(ms_img) / 2
        if out is not None:
            out["image"] = {"metric": "pixel*freq image solves/s (Solver::image_feautrier_order_2_new_imager)", "value": npix * args.image_freqs / (ms_img * 1e-3),
                            "unit": "solves/s", "ms_per_image": ms_img, "pixels": [nx, ny], "nfreqs": args.image_freqs,
                            "kernel_ms": {"trace": ist.ms_trace, "solve": ist.ms_solve}, "entries_per_pixel": ist.n_entries / max(1, ist.n_pixels),
                            "note": "end to end through mgb_image_feautrier_order_2 with host buffers (D2H of the image inside the timed region); "
                                    f"{flat['npoints']}-point model of this bench (the 10M-point cloud of configs[4] is not generated here); pixels/{world} per rank"}
    ctx.close()
    if out is not None and world == 1 and not args.no_levelpop:
        out["s_per_levelpop_iter"] = levelpop_iteration(args, flat, local)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        emit(out)


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
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
