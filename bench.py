#!/usr/bin/env python
"""bench.py — headline benchmark of the spgx hot path (contract: see DESIGN.md, "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl spgx|reference] [--config 3|5|4|2|1]

A "step" is one Solver::step() (one frame, dt = 0.01) on the workload:
  N = 1 : BASELINE config 3 — cloth 1024x1024 (1 048 576 vertices), StVK membrane + discrete bending + 2 corner
          anchors, VBD 10 iterations/substep, 1 substep.
  N > 1 : the same cloth with 1024 rows PER GPU ((1024 N) x 1024 vertices), row-strip domain decomposition with a
          per-colour halo exchange (weak scaling); `--config 5` runs the fixed 4096x4096 cloth instead (strong scaling).
metric = VBD vertex-iterations per second (vertices x iterations x substeps x frames / time), whole job.
`value` is device-resident (state already in HBM), `e2e` goes through the public API with pinned host buffers
(upload x, v -> step -> download x, v every frame).  `--impl reference` times the reference's own OpenMP CPU path
(oracle/_ref, else the CPU port) on the host cores on a bounded sample (BASELINE config 1: the same scene at 64x64).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
for p in (str(ROOT), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

DT = 0.01
CONFIGS = {
    # name: (builder, args, solver, substeps, iterations)
    1: ("cloth", (64, 64), "vbd", 1, 10),
    2: ("cloth_xpbd", (256, 256), "xpbd", 20, 1),
    3: ("cloth", (1024, 1024), "vbd", 1, 10),
    4: ("beam", (149, 38, 38), "vbd", 1, 20),
    5: ("cloth", (4096, 4096), "vbd", 1, 10),
    # SURVEY 8f rank 2 (not a BASELINE config): ImplicitEulerNewtonDx, one linearised solve per substep
    6: ("cloth", (1024, 1024), "newton_dx", 1, 1),
}


def build_scene(builder, args):
    from spg_b200 import scenes as S
    if builder == "cloth":
        return S.cloth(*args)
    if builder == "cloth_xpbd":
        return S.cloth(*args, S.K_SPRING, S.K_QUADRATIC_BENDING)
    if builder == "beam":
        return S.tet_beam(*args, S.K_TET_SNH)
    raise ValueError(builder)


def workload_name(builder, args, solver, substeps, iterations):
    if solver == "newton_dx":
        return (f"cloth {args[0]}x{args[1]} StVK membrane + discrete bending + 2 anchors, ImplicitEulerNewtonDx, {substeps} substep, "
                f"dt {DT}, CG tolerance 1e-5, <= 300 iterations")
    if builder == "cloth":
        return (f"cloth {args[0]}x{args[1]} StVK membrane + discrete bending + 2 anchors, VBD {iterations} iters/substep, "
                f"{substeps} substep, dt {DT}")
    if builder == "cloth_xpbd":
        return (f"cloth {args[0]}x{args[1]} springs + quadratic bending + 2 anchors, XPBD {substeps} substeps x "
                f"{iterations} pass, dt {DT}")
    return (f"Stable Neo-Hookean tet beam {args[0]}x{args[1]}x{args[2]} + anchors, VBD {iterations} iters/substep, "
            f"{substeps} substep, dt {DT}")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe).  Only samples taken
    UNDER LOAD count (GPU utilisation above 50 %, or power at least 100 W above the idle draw seen first): a 60 ms timed
    region inside a one-second sampling window used to report the idle clock (120 MHz) as the median.  `hold()` keeps the
    same workload running until enough loaded samples exist when the timed region itself is too short to be sampled."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,utilization.gpu,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device=0, period_ms=50):
        self.device = device
        self.period_ms = period_ms
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", str(self.period_ms), "-i", str(self.device)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def _parse(self):
        rows = []
        for ln in list(self.lines):
            f = [a.strip() for a in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                rows.append((float(f[1]), float(f[2]), float(f[3]), float(f[4]), f[5:9]))
            except ValueError:
                continue
        return rows

    @staticmethod
    def _loaded(rows):
        if not rows:
            return []
        idle_w = min(r[2] for r in rows)
        return [r for r in rows if r[3] > 50.0 or r[2] > idle_w + 100.0]

    def hold(self, work, min_samples=8, max_seconds=3.0):
        """keep calling work() (the timed workload, results discarded) until min_samples loaded samples exist"""
        if not self.proc:
            return
        t0 = time.perf_counter()
        while len(self._loaded(self._parse())) < min_samples and time.perf_counter() - t0 < max_seconds:
            work()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        rows = self._parse()
        loaded = self._loaded(rows)
        reasons = set()
        for r in loaded:
            for name, val in zip(self.NAMES, r[4]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median([r[0] for r in loaded])) if loaded else None,
                "sm_max_mhz": rows[-1][1] if rows else None, "reasons": sorted(reasons),
                "power_w_max": max(r[2] for r in loaded) if loaded else None,
                "samples": len(rows), "samples_under_load": len(loaded),
                "note": "median over the samples taken under load (utilisation > 50 % or power > idle + 100 W)"}


def peaks():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def time_oracle(cfg_id, frames, warmup, threads=None, min_seconds=0.0):
    """The reference's OpenMP path (or the CPU port) on the host cores: returns (units/s, ms/frame, info).
    min_seconds > 0: keep stepping (whole frames) until about that much CPU time has been measured."""
    import oracle_api
    builder, args, solver, substeps, iterations = CONFIGS[cfg_id]
    oracle = oracle_api.load()
    # all the host threads this process may use (torchrun exports OMP_NUM_THREADS=1 to its workers: override it)
    threads = threads or len(os.sched_getaffinity(0))
    oracle.set_threads(threads)
    cores = threads
    sc = build_scene(builder, args)
    og = oracle_api.OracleGroup(oracle, sc)
    osol = oracle_api.OracleSolver(og, {"vbd": 0, "xpbd": 1, "newton_dx": 4}[solver], DT, substeps, iterations)
    osol.step(max(warmup, 1))  # first frame includes the (quadratic) colouring
    t = time.perf_counter()
    osol.step(frames)
    el = time.perf_counter() - t
    cg_total = int(osol.cg_iterations().sum()) if solver == "newton_dx" else 0
    if el < min_seconds:  # a sample of about min_seconds of CPU work (the same work every frame: no adaptive counts)
        extra = min(int((min_seconds - el) / (el / frames)) + 1, 400)
        t = time.perf_counter()
        osol.step(extra)
        el += time.perf_counter() - t
        frames += extra
        cg_total += int(osol.cg_iterations().sum()) if solver == "newton_dx" else 0
    units = sc.n_vertices * cg_total if solver == "newton_dx" else \
        sc.n_vertices * iterations * substeps * frames if solver == "vbd" else \
        sum(e.stencils.shape[0] for e in sc.energies) * iterations * substeps * frames
    return units / el, 1e3 * el / frames, {"kind": oracle.kind, "cores": cores,
                                            "sample": f"{workload_name(*CONFIGS[cfg_id])}, {frames} frames after "
                                                      f"{max(warmup, 1)} warm-up"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    builder, a, solver, substeps, iterations = CONFIGS[args.config]
    # bounded sample of the workload: same scene family at the reference's own CPU-runnable size (config 1 / 2 at 64^2)
    sample_cfg = 1 if solver == "vbd" and builder == "cloth" else args.config
    if solver == "newton_dx":
        CONFIGS[60] = ("cloth", (64, 64), "newton_dx", substeps, iterations)
        sample_cfg = 60
    if builder == "beam":
        CONFIGS[40] = ("beam", (17, 5, 5), "vbd", substeps, iterations)
        sample_cfg = 40
    if builder == "cloth_xpbd":
        CONFIGS[20] = ("cloth_xpbd", (64, 64), "xpbd", substeps, iterations)
        sample_cfg = 20
    value, ms, info = time_oracle(sample_cfg, args.steps, args.warmup)
    unit = "vertex-iters/s" if solver == "vbd" else ("vertex-cg-iters/s" if solver == "newton_dx" else "constraint-projections/s")
    line = {"impl": "reference", "metric": metric_name(solver), "value": value, "unit": unit, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(builder, a, solver, substeps, iterations),
                       "note": "CPU arm timed on a bounded sample of the workload; the reference's rate is size-independent: 1.82e4 / 1.95e4 / 2.10e4 "
                               "vertex-iterations/s at 64^2 / 256^2 / 1024^2 on the same 8 cores (profiles/r2p_cpu_reference_rate.txt, "
                               "tools/cpu_rate_at_size.py)"},
            "cpu_baseline": {"value": value, "unit": unit, **info},
            "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def metric_name(solver):
    if solver == "newton_dx":
        return "implicit_euler_newton_dx_cg_vertex_iters_per_sec"
    return "vbd_vertex_iters_per_sec" if solver == "vbd" else "xpbd_constraint_projections_per_sec"


def run_implicit(args):
    """config 6: ImplicitEulerNewtonDx (implicitEulerNewtonDx.cpp:11-64) on the config-3 cloth.  A "step" is one frame =
    system force + Hessian assembly + one Jacobi-PCG solve (tolerance 1e-5, <= 300 iterations, the reference's constants).
    metric = vertices x CG iterations per second (the CG iteration is the unit of work: its count is data dependent and
    equal to the reference's); the roofline is quoted on the CG iteration (block-sparse product + two vector passes)."""
    from spg_b200 import api
    builder, a, solver, substeps, iterations = CONFIGS[args.config]
    if args.rows:
        a = (args.rows, a[1])
    sc = build_scene(builder, a)
    ctx = api.Context(int(os.environ.get("LOCAL_RANK", "0")))
    g = api.Group.from_scene(ctx, sc)
    g.finalize(api.SPGX_NEWTON)
    V = sc.n_vertices
    # every frame starts from the scene's initial state (flat cloth at rest, falling), so that every timed frame is the same
    # work: the reference's formulation (no PSD projection) hits its cap of 300 CG iterations from the first frame on a fine
    # cloth, unconverged solves compound over consecutive frames, and the device path follows the reference frame by frame in
    # that regime (tests/test_implicit.py::test_gpu_newton_dx_consecutive_frames_follow_the_reference, pinned at 96 x 96)
    class T:  # accumulated spgx_timings
        total_ms = solve_ms = inertial_ms = update_ms = 0.0
        kernel_launches = 0
    t = T()
    cg_list = []

    def frame(timed):
        g.reset()
        g.implicit_step(api.SPGX_IE_NEWTON_DX, DT, substeps, frames=1)
        if timed:
            tt = g.timings()
            for k in ("total_ms", "solve_ms", "inertial_ms", "update_ms", "kernel_launches"):
                setattr(t, k, getattr(t, k) + getattr(tt, k))
            cg_list.extend(g.cg_iterations().tolist())

    for _ in range(max(args.warmup, 3)):
        frame(False)
    sampler = ClockSampler(ctx.device)
    sampler.start()
    time.sleep(0.3)
    for _ in range(args.steps):
        frame(True)
    cg = np.array(cg_list)
    time.sleep(0.2)
    clocks = sampler.stop()
    rowptr, cols, _ = g.system_stiffness()
    nnzb = int(cols.size)
    cg_total = int(cg.sum())
    # per CG iteration: blocks (72 B) + column ids (4 B) + row bounds, p gathered once per row + 13 vector passes of 24 B
    # per vertex (product: p in, tmp out; update: p, tmp, r, invdiag, x in, x, r, z out; direction: z, p in, p out)
    bytes_cg = nnzb * 76.0 + (V + 1) * 4.0 + 13 * 24.0 * V
    peak, peak_src = peaks()
    achieved = bytes_cg * cg_total / (t.solve_ms * 1e-3) / 1e9
    value = V * cg_total / (t.total_ms * 1e-3)
    xh, vh = api.pinned_empty((V, 3)), api.pinned_empty((V, 3))
    x0h, v0h = api.pinned_empty((V, 3)), api.pinned_empty((V, 3))
    g.reset()
    g.download_into(x0h, v0h)
    ef = max(min(args.steps, 5), 1)
    t0 = time.perf_counter()
    cg_e2e = 0
    for _ in range(ef):
        g.upload(x0h, v0h)
        g.implicit_step(api.SPGX_IE_NEWTON_DX, DT, substeps, frames=1)
        cg_e2e += int(g.cg_iterations().sum())
        g.download_into(xh, vh)
    e2e_s = time.perf_counter() - t0
    assert np.isfinite(xh).all()
    cpu = None
    if not args.no_cpu_baseline:
        import oracle_api
        oracle = oracle_api.load()
        oracle.set_threads(len(os.sched_getaffinity(0)))
        scs = build_scene("cloth", (64, 64))
        osol = oracle_api.OracleSolver(oracle_api.OracleGroup(oracle, scs), 4, DT, 1, 1)
        osol.step(1)
        t1 = time.perf_counter()
        osol.step(3)
        el = time.perf_counter() - t1
        cpu = {"value": scs.n_vertices * int(osol.cg_iterations().sum()) / el, "unit": "vertex-cg-iters/s",
               "ms_per_frame_on_sample": 1e3 * el / 3, "kind": oracle.kind, "cores": oracle.max_threads(),
               "sample": "cloth 64x64, ImplicitEulerNewtonDx, 3 frames after 1 warm-up"}
    line = {"metric": metric_name(solver), "value": value, "unit": "vertex-cg-iters/s", "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": t.total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"cloth {a[0]}x{a[1]} StVK membrane + discrete bending + 2 anchors, ImplicitEulerNewtonDx, "
                                   f"{substeps} substep, dt {DT}, CG tolerance 1e-5, <= 300 iterations",
                       "vertices": V, "blocks_3x3": nnzb, "cg_iterations_per_frame": cg_total / args.steps,
                       "l2": "inputs larger than L2 (0.9 GB of matrix blocks per CG iteration vs 126 MB L2)",
                       "launch": "plain launches, the host polls the convergence flag every 8 iterations",
                       "state": "every frame starts from the scene's initial state", "parallelism": "1 GPU"},
            "clocks": clocks,
            "phases_ms_per_step": {"force_and_hessian_assembly": t.inertial_ms / args.steps, "cg_solve": t.solve_ms / args.steps,
                                   "update": t.update_ms / args.steps},
            "e2e": {"value": V * cg_e2e / e2e_s, "unit": "vertex-cg-iters/s", "h2d_bytes_per_step": int(xh.nbytes + vh.nbytes),
                    "d2h_bytes_per_step": int(xh.nbytes + vh.nbytes), "ms_per_step": 1e3 * e2e_s / ef, "steps": ef},
            "gpu_launches": int(t.kernel_launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "kernel": "bsrSpmvKernel + cgUpdateKernel + cgDirectionKernel (one CG iteration)",
                         "launches_per_step": 3 * cg_total / args.steps, "avg_launch_us": 1e3 * t.solve_ms / max(3 * cg_total, 1),
                         "algorithmic_bytes_per_launch": bytes_cg / 3, "algorithmic_bytes_per_unit": bytes_cg / V,
                         "kernel_share_of_step": t.solve_ms / t.total_ms, "peak_source": peak_src},
            "cpu_baseline": cpu}
    traffic, _, tsrc = ncu_traffic(args.config)
    line["roofline"]["traffic"] = traffic
    line["roofline"]["traffic_source"] = tsrc
    line["roofline"]["kernel"] = "bsrSpmvPipelinedKernel + cgUpdateKernel + cgDirectionKernel (one CG iteration)"
    print(json.dumps(line), flush=True)
    return 0


# the sources a config's dominant kernel is compiled from (spg_b200/csrc): an ncu capture of that kernel stays valid while
# THESE are unchanged; None = every source of the library
KERNEL_SOURCES = {
    3: ["vbd_stream.cuh", "vbd_stream_cloth_wedge.cu", "vbd_stream_cloth.cu", "streams.cuh", "energies.cuh", "dual.cuh", "linalg.cuh",
        "halo.cuh", "kernels.h", "kinds.def"],
    6: ["newton.cu", "hinge_hessian.cuh", "membrane_hessian.cuh", "energies.cuh", "dual.cuh", "linalg.cuh", "kernels.h", "kinds.def"],
}


def kernel_source_hash(cfg_id=None):
    """sha256 over the CUDA sources the config's dominant kernel is built from (all sources of the library without a config):
    ties an ncu capture (profiles/traffic.json) to a build of that kernel"""
    import hashlib
    h = hashlib.sha256()
    names = KERNEL_SOURCES.get(cfg_id)
    files = sorted((ROOT / "spg_b200" / "csrc").glob("*")) if names is None else [ROOT / "spg_b200" / "csrc" / n for n in sorted(names)]
    for f in files:
        if f.suffix in (".cu", ".cuh", ".h", ".def"):
            h.update(f.name.encode())
            h.update(f.read_bytes())
    return h.hexdigest()[:16]


def ncu_traffic(cfg_id):
    """(traffic bytes per launch, FP64 pipe %, provenance) of the dominant kernel from the committed ncu capture -- only
    when that capture was taken on THIS build of the kernels (same source hash); otherwise null with the reason."""
    prof = ROOT / "profiles" / "traffic.json"
    here = kernel_source_hash(cfg_id)
    if not prof.exists():
        return None, None, {"csrc_sha256": here, "note": "no ncu capture committed"}
    try:
        tj = json.loads(prof.read_text())
    except Exception as e:
        return None, None, {"csrc_sha256": here, "note": f"profiles/traffic.json unreadable: {e}"}
    there = tj.get(f"_csrc_sha256_{cfg_id}")
    src = {"csrc_sha256": here, "ncu_csrc_sha256": there, "hashed_sources": KERNEL_SOURCES.get(cfg_id, "all of spg_b200/csrc"),
           "ncu_capture": tj.get(f"_source_{cfg_id}", tj.get("_source"))}
    if there != here:
        src["note"] = "the committed ncu capture is of another build of this kernel: traffic omitted"
        return None, None, src
    return tj.get(str(cfg_id)), tj.get(f"fp64_pipe_pct_{cfg_id}"), src


def measure_e2e(api, ctx, g, step, x_init, units_per_frame, unit, steps, lo=0, hi=None, olo=0, ohi=None, sync_all=None):
    """End to end through the C ABI with pinned host buffers: every frame uploads its inputs (x, v), steps, and reads its
    result (x, v) back.  Headline = the pipelined frame loop (spgx_upload_async / spgx_vbd_step(sync = 0) /
    spgx_download_async): frames take their inputs from the host independently of each other (a frame queue), so the
    upload of frame k + 1 and the read-back of frame k - 1 run on their own streams while frame k is solved.  The serial
    loop (blocking upload -> step -> blocking download) is reported beside it.  Every frame's input is the scene's initial
    state held in pinned host memory (x_init, zero velocities), so every frame does the same work and returns the same
    result (checked).  [lo, hi) / [olo, ohi): index ranges a partitioned rank uploads / downloads."""
    V = x_init.shape[0]
    hi = V if hi is None else hi
    ohi = V if ohi is None else ohi
    xin, vin = api.pinned_empty((V, 3)), api.pinned_empty((V, 3))
    xout = [api.pinned_empty((V, 3)), api.pinned_empty((V, 3))]
    vout = [api.pinned_empty((V, 3)), api.pinned_empty((V, 3))]
    xin[lo:hi] = x_init[lo:hi]
    vin[lo:hi] = 0.0
    ef = max(min(steps, 200), 1)   # the K frames the caller asked for (the fill and drain of the frame queue -- the first upload, the last read-back -- are inside the timed region)
    sync_all = sync_all or (lambda: None)

    def serial():
        g.upload(xin, vin)
        step(1)
        g.download_into(xout[0], vout[0])
    for _ in range(2):
        serial()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(ef):
        serial()
    sync_all()
    serial_s = time.perf_counter() - t0
    assert np.isfinite(xout[0][olo:ohi]).all()
    x_serial = xout[0][olo:ohi].copy()

    # pipelined: inputs of every frame come from pinned host memory, results go to alternating pinned buffers
    def pipelined(n):
        for k in range(n):
            g.upload_async(xin, vin)
            step(1, sync=False)
            g.download_async(xout[k & 1], vout[k & 1])
        ctx.sync()
    pipelined(2)
    sync_all()
    t0 = time.perf_counter()
    pipelined(ef)
    sync_all()
    pipe_s = time.perf_counter() - t0
    # the pipelined loop returns, frame after frame, what the serial loop returns (same inputs, same arithmetic)
    assert np.array_equal(xout[(ef - 1) & 1][olo:ohi], x_serial) and np.array_equal(xout[ef & 1][olo:ohi], x_serial)
    nb_in, nb_out = 2 * 24 * (hi - lo), 2 * 24 * (ohi - olo)
    return {"value": units_per_frame * ef / pipe_s, "unit": unit, "h2d_bytes_per_step": int(nb_in), "d2h_bytes_per_step": int(nb_out),
            "ms_per_step": 1e3 * pipe_s / ef, "steps": ef,
            "pipeline": "frame queue: H2D of frame k+1 and D2H of frame k-1 on their own streams while frame k is solved "
                        "(spgx_upload_async / step(sync=0) / spgx_download_async); every frame's x, v come from and return to pinned host memory",
            "serial_ms_per_step": 1e3 * serial_s / ef, "serial_value": units_per_frame * ef / serial_s}


def measure_single(ctx, cfg_id, steps, warmup, rows=0, e2e=True, cpu=True, clocks=True):
    """One config on one GPU: device-resident value (CUDA-graph replays, CUDA events on the library's stream), the
    colour-pass / projection-pass roofline (plain launches, per-phase events), end to end through the C ABI, CPU baseline."""
    from spg_b200 import api
    builder, a, solver, substeps, iterations = CONFIGS[cfg_id]
    if rows and builder.startswith("cloth"):
        a = (rows, a[1])
    t0 = time.perf_counter()
    sc = build_scene(builder, a)
    scene_s = time.perf_counter() - t0
    g = api.Group.from_scene(ctx, sc)
    t0 = time.perf_counter()
    g.finalize(api.SPGX_VBD if solver == "vbd" else api.SPGX_XPBD)
    finalize_s = time.perf_counter() - t0   # host colouring + incidence + stream build + upload (blocking call)
    V = sc.n_vertices
    n_stencils = sum(e.stencils.shape[0] for e in sc.energies)

    def step(frames, sync=True):
        if solver == "vbd":
            g.vbd_step(DT, substeps, iterations, frames=frames, sync=sync)
        else:
            g.xpbd_step(DT, substeps, iterations, frames=frames, sync=sync)

    units_per_frame = V * iterations * substeps if solver == "vbd" else n_stencils * iterations * substeps
    unit = "vertex-iters/s" if solver == "vbd" else "constraint-projections/s"
    bytes_iter, bytes_lower = g.model_bytes()  # per VBD iteration (all colours) / per XPBD substep

    # ---- device-resident value: CUDA-graph replay of K frames, timed with CUDA events on the library's stream ----
    step(max(warmup, 3))
    sampler = ClockSampler(ctx.device) if clocks else None
    if sampler:
        sampler.start()
        time.sleep(0.15)
    step(steps)
    t = g.timings()
    total_ms = t.total_ms
    launches = int(t.kernel_launches)
    clk = None
    if sampler:
        sampler.hold(lambda: step(steps))   # same workload, results discarded: enough loaded samples for a median
        clk = sampler.stop()
    value = units_per_frame * steps / (total_ms * 1e-3)

    # ---- dominant kernel (the colour pass): plain launches with per-phase events over the same region ----
    g.set_graph_mode(False)
    step(1)
    rf = max(min(steps, 5), 1)
    step(rf)
    tp = g.timings()
    g.set_graph_mode(True)
    ncol = g.num_colors() if solver == "vbd" else None
    passes_per_frame = substeps * iterations * (ncol if solver == "vbd" else sum(g.num_colors(e) for e in range(len(sc.energies))))
    solve_ms_per_frame = tp.solve_ms / rf
    sweeps = iterations * substeps if solver == "vbd" else substeps
    info = g.kernel_info()
    achieved = bytes_iter * sweeps / (solve_ms_per_frame * 1e-3) / 1e9
    peak, peak_src = peaks()
    kernel = ("vbdStreamKernel" if g.streamed() else "vbdPassKernel") if solver == "vbd" else "xpbdPassKernel"
    traffic, fp64, tsrc = ncu_traffic(cfg_id)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": tsrc, "kernel": kernel,
                "launches_per_step": passes_per_frame,
                "avg_launch_us": 1e3 * solve_ms_per_frame / max(passes_per_frame, 1),
                "algorithmic_bytes_per_launch": bytes_iter * sweeps / max(passes_per_frame, 1),
                "algorithmic_bytes_per_unit": bytes_iter * sweeps / units_per_frame,
                "lower_bound_frac": bytes_lower * sweeps / (solve_ms_per_frame * 1e-3) / 1e9 / peak,
                "kernel_share_of_step": solve_ms_per_frame / (tp.total_ms / rf), "peak_source": peak_src}
    if solver == "vbd":
        # variant: resident blocks + 10 x "records through the shared-memory ring"; wedge_records: fan triangles whose membrane and
        # two hinges are ONE record of the colour pass (streams.cuh: ST_WEDGE)
        roofline["kernel_variant"] = {"lanes": info["lanes"], "variant": info["variant"], "wedge_records": info["wedge_records"]}
    # the same bytes over the whole step AS TIMED (graph replay, predict / finalize included): `frac` above is measured with
    # per-phase events in plain-launch mode, where a launch-bound sequence (XPBD: 540 passes of a few microseconds) pays the
    # full launch latency per pass that the replayed graph does not
    roofline["frac_of_timed_step"] = bytes_iter * sweeps / (total_ms / steps * 1e-3) / 1e9 / peak
    if fp64 is not None:
        roofline["fp64_pipe_pct_ncu"] = fp64   # the FP64 pipe co-binds this kernel
    if traffic:
        roofline["dram_frac_of_peak"] = traffic / (roofline["avg_launch_us"] * 1e-6) / 1e9 / peak

    out = {"value": value, "unit": unit, "ms_per_step": total_ms / steps, "launches": launches, "clocks": clk,
           "roofline": roofline, "vertices": V, "stencils": n_stencils, "colours": ncol, "finalize_s": finalize_s,
           "scene_build_s": scene_s, "workload": workload_name(builder, a, solver, substeps, iterations), "solver": solver}

    # ---- end to end through the public API: pinned host buffers, H2D + step + D2H every frame ----
    if e2e:
        out["e2e"] = measure_e2e(api, ctx, g, step, sc.x, units_per_frame, unit, steps)

    # ---- CPU baseline on the host cores, bounded sample ----
    if cpu:
        try:
            sample_cfg = 1 if (solver == "vbd" and builder == "cloth") else None
            if builder == "beam":
                CONFIGS[40] = ("beam", (17, 5, 5), "vbd", substeps, iterations)
                sample_cfg = 40
            if builder == "cloth_xpbd":
                CONFIGS[20] = ("cloth_xpbd", (64, 64), "xpbd", substeps, iterations)
                sample_cfg = 20
            v_cpu, ms_cpu, cinfo = time_oracle(sample_cfg, 3, 1, min_seconds=10.0)
            out["cpu_baseline"] = {"value": v_cpu, "unit": unit, "ms_per_frame_on_sample": ms_cpu, **cinfo}
        except Exception as e:  # the oracle is test infrastructure; its absence must not hide the GPU number
            out["cpu_baseline"] = {"value": None, "unit": unit, "error": str(e)}
    g.close()
    return out


def run_single(args):
    from spg_b200 import api
    ctx = api.Context(int(os.environ.get("LOCAL_RANK", "0")))
    m = measure_single(ctx, args.config, args.steps, args.warmup, rows=args.rows, cpu=not args.no_cpu_baseline)
    solver = m["solver"]
    line = {"metric": metric_name(solver), "value": m["value"], "unit": m["unit"], "n_gpus": 1, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": m["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": m["workload"], "vertices": m["vertices"], "stencils": m["stencils"], "colours": m["colours"],
                       "l2": "inputs larger than L2 (element records + incidence ~0.4 GB per iteration vs 126 MB L2)",
                       "launch": "CUDA graph replay of one step()", "parallelism": "1 GPU"},
            "clocks": m["clocks"], "e2e": m.get("e2e"), "gpu_launches": m["launches"], "roofline": m["roofline"],
            "cpu_baseline": m.get("cpu_baseline"), "finalize_s": m["finalize_s"], "scene_build_s": m["scene_build_s"]}
    # the other single-GPU BASELINE configs ride along (they cost milliseconds of device time): config 2 (XPBD cloth
    # 256 x 256) and config 4 (stable Neo-Hookean beam, ~1 M tets) -- parity-gated at full size by tests/test_gpu_atsize.py
    if args.config == 3 and not args.rows and not args.no_other_configs:
        others = {}
        for cid in (2, 4):
            try:
                o = measure_single(ctx, cid, min(args.steps, 10), args.warmup, e2e=True, cpu=False, clocks=False)
                others[f"config{cid}"] = {k: o[k] for k in ("workload", "value", "unit", "ms_per_step", "launches", "vertices", "stencils",
                                                             "colours", "finalize_s", "e2e")}
                others[f"config{cid}"]["roofline"] = {k: o["roofline"][k] for k in ("frac", "achieved", "peak", "kernel", "launches_per_step",
                                                                                    "avg_launch_us", "algorithmic_bytes_per_unit",
                                                                                    "lower_bound_frac", "kernel_share_of_step")}
            except Exception as e:
                others[f"config{cid}"] = {"error": str(e)}
        line["other_configs"] = others
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="spgx", choices=["spgx", "reference"])
    ap.add_argument("--config", type=int, default=3, choices=sorted(CONFIGS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="N = 1: skip the config 2 / config 4 blocks")
    ap.add_argument("--rows", type=int, default=0, help="diagnostics: override the first grid dimension of a cloth config")
    ap.add_argument("--no-identity", action="store_true", help="N > 1: skip the N-GPU == 1-GPU bit-identity check")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the config-5 strong-scaling block")
    ap.add_argument("--halo", default="p2p", choices=["p2p", "nccl"], help="multi-GPU halo exchange (bench_multi.py)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 or world > 1:
        from bench_multi import run_multi
        return run_multi(args)
    if CONFIGS[args.config][2] == "newton_dx":
        return run_implicit(args)
    return run_single(args)


if __name__ == "__main__":
    sys.exit(main())
