"""Multi-GPU arm of bench.py: one process per GPU (torchrun), torch.distributed (NCCL) for the plumbing -- rendezvous,
the broadcast of the NCCL unique id the library's own communicator is built from, barriers and the max-over-ranks
reduction of the device times.  The data path (per-colour halo exchange inside the VBD sweep) is the library's own
fused peer-memory exchange (spg_b200/csrc/halo.cuh: boundary vertices stored into the neighbours' halo slots by the
colour-pass kernel itself, event counters in peer memory; `--halo nccl` selects the ncclSend / ncclRecv schedule
captured in the step graph instead, spg_b200/csrc/comm.cu).

Workloads (row-strip domain decomposition, global colouring):
  default / --config 3 : the headline line is WEAK scaling, cloth (1024 N) x 1024 (1024 rows per GPU), VBD 10 iterations
                         (the driver's efficiency formula applies to it); the same line carries a `strong_config5` block:
                         BASELINE config 5 (cloth 4096 x 4096) on the N GPUs, its time on ONE GPU measured in the same job
                         on rank 0, the speed-up, and `identity_vs_1gpu` (below)
  --config 5           : strong scaling only, cloth 4096 x 4096 on N GPUs
  --config 4           : strong scaling, stable Neo-Hookean beam 149 x 38 x 38 on N GPUs (slabs along x)
`identity_vs_1gpu`: after the timed region every rank restarts from the initial state, steps one frame and hashes the
positions and velocities of the vertices it owns; rank 0 steps the SAME scene as a single domain on its own GPU and hashes
the same index ranges: true iff all hashes agree bit for bit (the N-GPU result IS the 1-GPU result, which
tests/test_gpu_atsize.py ties to the reference).
The scene is built once (rank 0) and shared with the other ranks through /dev/shm; every rank still finalizes against
the global stencil lists (the colouring is global by definition: vbd.cpp:203-229).
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys
import time
from pathlib import Path

import numpy as np

import bench as B


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def shared_scene(builder, a, rank, dist, tag):
    """rank 0 builds the scene and leaves its arrays in /dev/shm; the others map them (no N-fold host work)."""
    from spg_b200 import scenes as S
    root = Path("/dev/shm") if Path("/dev/shm").is_dir() else Path("/tmp")
    d = root / f"spgx_scene_{os.environ.get('MASTER_PORT', '0')}_{tag}"
    t0 = time.perf_counter()
    if rank == 0:
        sc = B.build_scene(builder, a)
        shutil.rmtree(d, ignore_errors=True)
        d.mkdir(parents=True)
        np.save(d / "x.npy", sc.x)
        np.save(d / "x0.npy", sc.x0)
        np.save(d / "mass.npy", sc.mass)
        np.save(d / "kinds.npy", np.array([e.kind for e in sc.energies], dtype=np.int64))
        for i, e in enumerate(sc.energies):
            np.save(d / f"st{i}.npy", e.stencils)
            np.save(d / f"pr{i}.npy", e.params)
    dist.barrier()
    if rank != 0:
        sc = S.Scene(np.load(d / "x.npy", mmap_mode="r"), np.load(d / "x0.npy", mmap_mode="r"), np.load(d / "mass.npy", mmap_mode="r"))
        for i, k in enumerate(np.load(d / "kinds.npy")):
            sc.energies.append(S.SceneEnergy(int(k), np.load(d / f"st{i}.npy", mmap_mode="r"), np.load(d / f"pr{i}.npy", mmap_mode="r")))
    dist.barrier()
    if rank == 0:
        shutil.rmtree(d, ignore_errors=True)  # mapped pages stay valid for the ranks that hold them
    return sc, time.perf_counter() - t0


class Job:
    """one partitioned VBD workload on all ranks"""

    def __init__(self, args, ctx, dist, torch, rank, world, builder, a, solver, substeps, iterations, tag):
        from spg_b200 import api
        self.api, self.args, self.ctx, self.dist, self.torch, self.rank, self.world = api, args, ctx, dist, torch, rank, world
        self.builder, self.a, self.solver, self.substeps, self.iterations = builder, a, solver, substeps, iterations
        self.sc, self.scene_s = shared_scene(builder, a, rank, dist, tag)
        self.V = self.sc.n_vertices
        g = api.Group.from_scene(ctx, self.sc)
        g.set_partition(rank, world)
        t0 = time.perf_counter()
        g.finalize(api.SPGX_VBD)
        self.finalize_s = self.max_over_ranks(time.perf_counter() - t0)
        if getattr(args, "halo", "p2p") == "p2p":
            # CUDA IPC handles all-gathered with torch.distributed; from here on the colour-pass kernel stores boundary
            # vertices straight into the neighbours' halo slots and synchronises through event counters in peer memory
            try:
                g.connect_peers(dist)
            except api.SpgxError as e:  # raised on EVERY rank when any rank cannot connect: all keep the NCCL exchange
                if rank == 0:
                    print(f"peer-memory exchange unavailable ({e}); using ncclSend/ncclRecv", file=sys.stderr, flush=True)
        self.halo_name = {api.SPGX_HALO_NCCL: "ncclSend/ncclRecv per colour, captured in the step graph",
                          api.SPGX_HALO_P2P: "peer-memory stores from the colour-pass kernel + event counters over NVLink (no NCCL on the data path)",
                          api.SPGX_HALO_NONE: "none"}[g.halo_mode()]
        self.g = g
        self.info = g.partition_info()

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def step(self, frames, sync=True, g=None):
        (g or self.g).vbd_step(B.DT, self.substeps, self.iterations, frames=frames, sync=sync)

    def timed(self, steps, warmup, sample_clocks):
        """K CUDA-graph replays per rank, device time by CUDA events on the library's stream, MAX over ranks"""
        dist, torch = self.dist, self.torch
        self.step(max(warmup, 3))
        sampler = B.ClockSampler(self.ctx.device) if (sample_clocks and self.rank == 0) else None
        dist.barrier()
        torch.cuda.synchronize()
        if sampler:
            sampler.start()
            time.sleep(0.15)
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        self.step(steps)
        torch.cuda.synchronize()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        dev_ms = self.g.timings().total_ms
        launches = int(self.g.timings().kernel_launches)
        dist.barrier()
        total_ms = self.max_over_ranks(dev_ms)
        wall_ms = self.max_over_ranks(wall_ms)
        clocks = None
        if sample_clocks:
            # the timed region is tens of milliseconds: keep the SAME workload running on every rank (a partitioned step
            # needs all of them) for about a second so that the sampler sees rank 0's GPU under load
            for _ in range(max(1, int(1000.0 / max(total_ms, 1.0)))):
                self.step(steps)
            if sampler:
                clocks = sampler.stop()
            dist.barrier()
        return total_ms, wall_ms, launches, clocks

    def roofline(self, steps):
        g = self.g
        g.set_graph_mode(False)
        self.step(1)
        rf = max(min(steps, 5), 1)
        self.dist.barrier()
        self.step(rf)
        tp = g.timings()
        g.set_graph_mode(True)
        solve_ms = self.max_over_ranks(tp.solve_ms / rf)
        frame_ms = self.max_over_ranks(tp.total_ms / rf)
        sweeps = self.iterations * self.substeps
        bytes_iter, bytes_lower = g.model_bytes()  # owned share of this rank
        bytes_all, lower_all = self.sum_over_ranks(bytes_iter), self.sum_over_ranks(bytes_lower)
        peak, peak_src = B.peaks()
        achieved = bytes_all * sweeps / (solve_ms * 1e-3) / 1e9
        ncol = g.num_colors()
        world = self.world
        return {"bound": "hbm", "achieved": achieved, "peak": peak * world, "unit": "GB/s", "frac": achieved / (peak * world),
                "traffic": None, "kernel": "vbdStreamKernel (halo exchange inside the phase: " + self.halo_name + ")",
                "launches_per_step": sweeps * ncol, "avg_launch_us": 1e3 * solve_ms / max(sweeps * ncol, 1),
                "algorithmic_bytes_per_unit": bytes_all * sweeps / (self.V * sweeps),
                "lower_bound_frac": lower_all * sweeps / (solve_ms * 1e-3) / 1e9 / (peak * world),
                "kernel_share_of_step": solve_ms / frame_ms, "peak_source": peak_src + f" x {world} GPUs"}

    def e2e(self, steps):
        """bench.measure_e2e on every rank (upload of the local index range + step + download of the owned range per frame),
        MAX over ranks of the wall time"""
        lo, hi, olo, ohi = self.info["local_range"] + self.info["owned_range"]

        def sync_all():
            self.torch.cuda.synchronize()
            self.dist.barrier()
        units = self.V * self.iterations * self.substeps
        r = B.measure_e2e(self.api, self.ctx, self.g, self.step, self.sc.x, units, "vertex-iters/s", steps, lo, hi, olo, ohi, sync_all)
        pipe_ms = self.max_over_ranks(r["ms_per_step"])
        serial_ms = self.max_over_ranks(r["serial_ms_per_step"])
        r.update({"ms_per_step": pipe_ms, "value": units / (pipe_ms * 1e-3), "serial_ms_per_step": serial_ms,
                  "serial_value": units / (serial_ms * 1e-3), "h2d_bytes_per_step": int(self.sum_over_ranks(r["h2d_bytes_per_step"])),
                  "d2h_bytes_per_step": int(self.sum_over_ranks(r["d2h_bytes_per_step"]))})
        return r, True

    def identity_vs_single(self, steps=0, warmup=3):
        """N-GPU result == 1-GPU result, bit for bit (see the module docstring).  steps > 0: rank 0 also times the single
        domain (CUDA-graph replays, CUDA events) and returns its ms per step."""
        api, g, dist = self.api, self.g, self.dist
        olo, ohi = self.info["owned_range"]
        g.reset()
        self.step(1)
        x, v = g.download()
        mine = (olo, ohi, _sha(x[olo:ohi]), _sha(v[olo:ohi]))
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine)
        lanes = g.kernel_info()["lanes"]
        result = [None, None]
        if self.rank == 0:
            # same lanes-per-vertex choice as the strips (it fixes the summation order inside a vertex)
            os.environ["SPGX_VBD_LANES"] = str(lanes)
            try:
                g1 = api.Group.from_scene(self.ctx, self.sc)
                g1.finalize(api.SPGX_VBD)
                self.step(1, g=g1)
                x1, v1 = g1.download()
                ok = all(_sha(x1[lo:hi]) == hx and _sha(v1[lo:hi]) == hv for lo, hi, hx, hv in everyone)
                ok = ok and sorted((lo, hi) for lo, hi, _, _ in everyone)[0][0] == 0 and sum(hi - lo for lo, hi, _, _ in everyone) == self.V
                ms1 = None
                if steps > 0:
                    self.step(max(warmup, 3), g=g1)
                    self.step(steps, g=g1)
                    ms1 = g1.timings().total_ms / steps
                g1.close()
                result = [bool(ok), ms1]
            finally:
                os.environ.pop("SPGX_VBD_LANES", None)
        dist.broadcast_object_list(result, src=0)
        return result[0], result[1]

    def close(self):
        self.g.disconnect_peers(self.dist)
        self.g.close()
        self.dist.barrier()


def run_multi(args):
    import torch
    import torch.distributed as dist
    from spg_b200 import api

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    builder, a, solver, substeps, iterations = B.CONFIGS[args.config]
    scaling = "strong"
    if args.config == 3:
        a = (a[0] * world, a[1])  # 1024 rows per GPU
        scaling = "weak"
    if solver != "vbd":
        raise SystemExit("multi-GPU runs cover the VBD configs (3, 4, 5)")

    ctx = api.Context(local_rank)
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid.copy_(torch.from_numpy(api.comm_unique_id()))
    dist.broadcast(uid, 0)
    ctx.init_comm(rank, world, uid.cpu().numpy())

    job = Job(args, ctx, dist, torch, rank, world, builder, a, solver, substeps, iterations, "main")
    total_ms, wall_ms, launches, clocks = job.timed(args.steps, args.warmup, sample_clocks=True)
    V = job.V
    value = V * iterations * substeps * args.steps / (total_ms * 1e-3)
    roofline = job.roofline(args.steps)
    e2e, finite = job.e2e(args.steps)
    identity, _ = job.identity_vs_single() if not args.no_identity else (None, None)
    halo_total = job.sum_over_ranks(job.info["n_local"] - job.info["n_owned"])
    line = {"metric": B.metric_name(solver), "value": value, "unit": "vertex-iters/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": B.workload_name(builder, a, solver, substeps, iterations), "vertices": V,
                       "colours": job.g.num_colors(), "parallelism": f"domain decomposition, {world} contiguous index ranges "
                       f"(row strips / slabs), per-colour halo exchange: {job.halo_name}",
                       "halo_vertices_total": int(halo_total), "launch": "CUDA graph replay of one step() per rank",
                       "l2": "inputs larger than L2 (incidence streams ~0.7 GB per sweep per 1M vertices)"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": None,
            "wall_ms_per_step": wall_ms / args.steps, "finite": finite, "identity_vs_1gpu": identity,
            "finalize_s": job.finalize_s, "scene_build_s": job.scene_s}
    job.close()

    # ---- BASELINE config 5 on the same N GPUs (strong scaling; north_star: >= 6x at 8 GPUs) ----
    if args.config == 3 and not args.no_strong:
        b5, a5, s5, sub5, it5 = B.CONFIGS[5]
        k5 = max(min(args.steps, 10), 1)
        job5 = Job(args, ctx, dist, torch, rank, world, b5, a5, s5, sub5, it5, "c5")
        ms5, _, launches5, _ = job5.timed(k5, args.warmup, sample_clocks=False)
        roof5 = job5.roofline(k5)
        ident5, ms1 = job5.identity_vs_single(steps=k5, warmup=args.warmup)
        halo5 = job5.sum_over_ranks(job5.info["n_local"] - job5.info["n_owned"])
        line["strong_config5"] = {"workload": B.workload_name(b5, a5, s5, sub5, it5), "vertices": job5.V, "n_gpus": world,
                                  "steps": k5, "ms_per_step": ms5 / k5, "value": job5.V * it5 * sub5 * k5 / (ms5 * 1e-3),
                                  "unit": "vertex-iters/s", "ms_per_step_1gpu": ms1,
                                  "speedup_vs_1gpu": (ms1 / (ms5 / k5)) if ms1 else None, "frac": roof5["frac"],
                                  "avg_launch_us": roof5["avg_launch_us"], "halo_vertices_total": int(halo5),
                                  "identity_vs_1gpu": ident5, "gpu_launches": launches5, "finalize_s": job5.finalize_s,
                                  "scene_build_s": job5.scene_s,
                                  "note": "the 1-GPU time is measured in this job on rank 0's GPU (same kernels, single domain)"}
        job5.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0
