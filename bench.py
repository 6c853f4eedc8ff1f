#!/usr/bin/env python
"""bench.py — throughput of the perturbator hot path on B200.

One "step" = one planner pass (BASELINE config 2: Rexarm 5-DOF, 128^3 uint8 grid,
64 waypoints x 1024 samples per GPU, 100 optimiser iterations of
noise -> FK -> collision gathers -> smoothness -> argmin -> keep-best update).
`value` = waypoint-samples costed per second over all ranks, inputs resident in
HBM, timed with CUDA events on the ctx stream; `e2e` = the same metric through
the host-buffer C-ABI call doapp_plan (H2D of start/goal/initial trajectory,
D2H of the best trajectory inside the timed region).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from tests import synth  # noqa: E402

METRIC = "waypoint_samples_per_sec"
UNIT = "waypoint-samples/s"

# BASELINE configs (per-GPU shapes; samples are weak-scaled over ranks)
WORKLOADS = {
    "c2": dict(desc="C2: Rexarm 5-DOF, 128^3 u8 grid (1 cm), 64 waypoints x 1024 samples/GPU, "
                    "100 optimizer iterations per step",
               grid_n=128, res=0.01, n=64, m=1024, iters=100, sigma=0.05, fill=0.0),
    "c3": dict(desc="C3: dynamic obstacle, 128^3 u8 grid rebuilt every iteration from a 300k-point "
                    "cloud (clear + scatter), then one optimizer iteration of 64 waypoints x 1024 "
                    "samples; 20 iterations per step",
               grid_n=128, res=0.01, n=64, m=1024, iters=20, sigma=0.05, fill=0.0, points=300_000),
    "c4": dict(desc="C4: 128 waypoints x 131072 samples/GPU, 128^3 u8 grid, 4 iterations per step",
               grid_n=128, res=0.01, n=128, m=131072, iters=4, sigma=0.05, fill=0.0),
    "c4m": dict(desc="C4 (whole 1M-sample config on each GPU): 128 waypoints x 1048576 samples/GPU, "
                     "128^3 u8 grid, 2 iterations per step",
                grid_n=128, res=0.01, n=128, m=1048576, iters=2, sigma=0.05, fill=0.0),
    "c2u": dict(desc="C2 with the URDF-correct FK model (13 collision spheres per waypoint): 128^3 u8 "
                     "grid, 64 waypoints x 1024 samples/GPU, 100 optimizer iterations per step",
                grid_n=128, res=0.01, n=64, m=1024, iters=100, sigma=0.05, fill=0.0, fk_model=True),
    "c4u": dict(desc="C4 with the URDF-correct FK model (SURVEY 8f row 3): 13 collision spheres per "
                     "waypoint instead of 5 link points; 128 waypoints x 131072 samples/GPU, 128^3 u8 "
                     "grid, 4 iterations per step",
                grid_n=128, res=0.01, n=128, m=131072, iters=4, sigma=0.05, fill=0.0, fk_model=True),
    "c5": dict(desc="C5: 512^3 u8 grid (128 MiB, L2-spill), 128 waypoints x 16384 samples/GPU, "
                    "4 iterations per step",
               grid_n=512, res=0.0025, n=128, m=16384, iters=4, sigma=0.5, fill=0.05),
}
D = 5
LINKS = 5


FK_MODEL_SPHERES = 13  # doapp_fk_model_rexarm


def alg_bytes_per_wps(n: int, cell_bytes: int = 1, links: int = LINKS) -> float:
    """5 gathers (13 with the FK model) x cell bytes + the sample's noise vector and
    its score/collision outputs amortised over its n waypoints."""
    return links * cell_bytes + (4 * D + 8) / n


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
                     0x4: "sw_power_cap", 0x80: "hw_power_brake"}
            while not self._halt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else \
                    nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
                time.sleep(0.02)
        except Exception as e:  # NVML missing: report that, do not fail the bench
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def make_inputs(w):
    grid = synth.box_grid(w["grid_n"], w["res"], fill=w["fill"])
    start, goal = synth.START, synth.GOALS[1]
    return grid, start, goal


def cpu_plan_rate(w, iters, threads, reps=1, min_seconds=0.0):
    """The oracle's planner loop (CPU restatement of the reference arithmetic) on the same
    workload, `iters` iterations, all host threads; repeated until `min_seconds` of CPU work have
    been timed; returns (wps/s of the best repetition, its seconds, sample text)."""
    import oracle
    oracle.build()
    grid, start, goal = make_inputs(w)
    orc = oracle.Oracle(k=1, m=w["m"], n=w["n"], d=D, sigma=w["sigma"], select_mode=1,
                        fk_model=oracle.fk_model_rexarm() if w.get("fk_model") else None)
    orc.set_grid(grid, w["res"])
    init = oracle.plan_lerp(start, goal, w["n"])
    best, total, done = float("inf"), 0.0, 0
    while done < reps or total < min_seconds:
        t0 = time.perf_counter()
        orc.plan(start, init, iters, None, threads=threads)
        dt = time.perf_counter() - t0
        best, total, done = min(best, dt), total + dt, done + 1
    wps = w["m"] * w["n"] * iters
    return wps / best, best, (f"{iters} iterations x {w['m']} samples x {w['n']} waypoints, best of "
                              f"{done} repetitions ({total:.1f} s of CPU work)")


def run_reference(args, w, rank):
    """--impl reference: the reference's CPU cost path.  The reference has no runnable CPU (or
    GPU) implementation of the full path (collision term stubbed), so this is the
    oracle port, threaded over samples the way cpu_distance_grid.cu:88-110 threads its fill."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    iters = w["iters"]
    w = dict(w, m=w["m"] * args.gpus)  # the whole job of the N-GPU arm (samples weak-scale)
    for _ in range(args.warmup):
        cpu_plan_rate(w, max(1, iters // 10), threads)
    secs = []
    for _ in range(args.steps):
        _, s, sample = cpu_plan_rate(w, iters, threads)
        secs.append(s)
    ms = 1e3 * float(np.mean(secs))
    val = w["m"] * w["n"] * iters / (ms * 1e-3)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": w["desc"], "samples_total": w["m"],
                       "note": "CPU oracle port on host cores, whole job of the N-GPU arm"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": sample + " per step"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# DRAM bytes (read + write) per cost-kernel launch from the committed `ncu --set full`
# captures (profiles/r01/ncu_full_cost_kernel_*_v3.raw.csv, dram__bytes_read/write.sum)
NCU_TRAFFIC = {"c2": 79.6e3, "c4": 84.2e3, "c5": 3.08e6}


def run_c3(args, w, rank, world, local_rank):
    """BASELINE config 3: every iteration rebuilds the grid from a fresh cloud (stage 5:
    cudaMemset clear + scatter_kernel), re-scores the kept trajectory and runs one optimizer
    iteration (doapp_replan_dynamic).  Replicas only at N > 1 (every rank rebuilds its own copy
    of the grid from the same cloud: cheaper than broadcasting it)."""
    import torch
    from doapp_b200 import capi
    torch.cuda.set_device(local_rank)
    m, n, iters, P = w["m"], w["n"], w["iters"], w["points"]
    grid, start, goal = make_inputs(w)
    init = capi.plan_lerp(start, goal, n)
    ctx = capi.Context(k=1, m=m, n=n, d=D, device=local_rank, sigma=w["sigma"], select_mode=1)
    ctx.set_grid(np.zeros_like(grid), w["res"])
    clouds = np.stack([synth.cloud(P, it, size=w["grid_n"] * w["res"]) for it in range(iters)])
    h_clouds = torch.from_numpy(clouds).pin_memory().numpy()
    for _ in range(max(3, args.warmup)):
        ctx.replan_dynamic(start, goal, init, h_clouds)
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = ctx.launch_count()
    dev_ms, t0 = 0.0, time.perf_counter()
    for _ in range(args.steps):
        r = ctx.replan_dynamic(start, goal, init, h_clouds)  # H2D of the clouds + loop + D2H
        dev_ms += r["elapsed_ms"]                             # CUDA events around the loop only
    e2e_s = time.perf_counter() - t0
    launches = ctx.launch_count() - l0
    clocks = sampler.stop()
    wps_step = m * n * iters
    ms_step = dev_ms / args.steps
    cells = w["grid_n"] ** 3
    alg = 12.0 * P + P + cells          # cloud read + byte stores + clear, per rebuild
    line = {
        "metric": METRIC, "value": wps_step / (ms_step * 1e-3), "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": w["desc"], "samples_per_gpu": m, "waypoints": n, "joints": D,
                   "iterations_per_step": iters, "points_per_cloud": P,
                   "grid": f"{w['grid_n']}^3 u8", "l2": "cloud of every iteration is new data "
                   "(72 MB per step streamed from HBM); grid L2-resident by design"},
        "iterations_per_sec": iters / (ms_step * 1e-3),
        "grid_rebuilds_per_sec": iters / (ms_step * 1e-3),
        "points_per_sec": P * iters / (ms_step * 1e-3),
        "e2e": {"value": wps_step * args.steps / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": int(h_clouds.nbytes + init.nbytes + start.nbytes),
                "d2h_bytes_per_step": int(init.nbytes + 8)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "scatter_kernel + clear", "achieved": None,
                     "peak": None, "unit": "GB/s", "frac": None, "traffic": None,
                     "alg_bytes_per_launch": alg,
                     "note": "latency-bound: 5.7 MB per rebuild is < 1 us of HBM time"},
        "clocks": clocks, "best_score": r["best_score"],
    }
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()


def run_ours(args, w, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from doapp_b200 import capi

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    m, n, iters = w["m"], w["n"], w["iters"]
    grid, start, goal = make_inputs(w)
    init = capi.plan_lerp(start, goal, n)
    ctx = capi.Context(k=1, m=m, n=n, d=D, device=local_rank, sigma=w["sigma"], select_mode=1,
                       sample_offset=rank * m, global_samples_per_climber=m * world)
    if world > 1:
        uid = torch.zeros(capi.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid = torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8).to(dev)
        dist.broadcast(uid, 0)
        ctx.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
        # the grid is broadcast once over NVLink (north_star); rank 0 holds the real cells
        if rank == 0:
            ctx.set_grid(grid, w["res"])
        else:
            ctx.grid_alloc(capi.GRID_U8, capi.FRAME_CDG, grid.shape[2], grid.shape[1],
                           grid.shape[0], w["res"])
        ctx.grid_broadcast(0)
        from doapp_b200 import sharding
        peer = sharding.connect_peers(ctx, rank, world, dev) if not os.environ.get("DOAPP_NO_PEER") else False
    else:
        ctx.set_grid(grid, w["res"])
        peer = False
    links = LINKS
    if w.get("fk_model"):
        ctx.set_fk_model(capi.fk_model_rexarm())
        links = FK_MODEL_SPHERES

    # pinned host staging for the e2e leg
    h_start = torch.from_numpy(start.copy()).pin_memory().numpy()
    h_goal = torch.from_numpy(goal.copy()).pin_memory().numpy()
    h_init = torch.from_numpy(init.copy()).pin_memory().numpy()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def flush_l2():
        flush.fill_(1)
        torch.cuda.synchronize()

    first = ctx.plan(h_start, h_goal, h_init, iters)  # stages inputs, builds the CUDA graph
    for _ in range(args.warmup):
        ctx.plan_bench(iters, 1)

    # ---- device-timed region: K steps, inputs resident in HBM -------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    l0 = ctx.launch_count()
    ms_total = 0.0
    for _ in range(args.steps):
        flush_l2()
        ms_total += ctx.plan_bench(iters, 1)
    launches = ctx.launch_count() - l0
    persistent = bool(ctx.plan_mode() & 1)
    peer = bool(ctx.plan_mode() & 2)
    barrier()
    ms_total = max_over_ranks(ms_total)
    clocks = sampler.stop()
    res = ctx.plan_wait()
    assert np.array_equal(res["best_traj"], first["best_traj"]), "bench result not reproducible"

    # ---- e2e: host buffers through doapp_plan, copies inside the timed region ---------
    for _ in range(args.warmup):
        ctx.plan(h_start, h_goal, h_init, iters)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = ctx.plan(h_start, h_goal, h_init, iters)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()

    # ---- dominant kernel: per-launch CUDA events around every cost launch -------------
    cost_ms, cost_launches, loop_ms = ctx.plan_profile(iters)
    cost_ms, cost_launches, loop_ms = ctx.plan_profile(iters)
    ctx.plan_wait()

    # ---- gather roofline: measured rate of random one-sector loads, L2- and HBM-resident ---
    cell_bytes = 1
    g_l2 = ctx.probe_gather(8 << 20, cell_bytes)
    g_hbm = ctx.probe_gather(1 << 30, cell_bytes)
    grid_bytes = w["grid_n"] ** 3 * cell_bytes
    g_peak = g_l2 if grid_bytes <= (64 << 20) else g_hbm

    wps_step = m * n * iters * world
    ms_step = ms_total / args.steps
    value = wps_step / (ms_step * 1e-3)
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    per_launch_s = cost_ms * 1e-3 / max(1, cost_launches)
    # one launch per iteration, or ONE persistent launch running all the iterations
    units_per_launch = m * n * iters / max(1, cost_launches)
    alg = alg_bytes_per_wps(n, links=links) * units_per_launch
    kernel_name = "plan_persistent_kernel" if cost_launches == 1 and iters > 1 else "cost_kernel"
    if w.get("fk_model"):
        kernel_name += "<FK model>"
    achieved = alg / per_launch_s / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["desc"], "samples_per_gpu": m, "waypoints": n, "joints": D,
                   "iterations_per_step": iters, "grid": f"{w['grid_n']}^3 u8",
                   "lookups_per_waypoint": links,
                   "parallelism": f"samples sharded over {world} GPU(s), grid replicated"
                                  + (", argmin keys merged in-kernel through NVLink peer mailboxes"
                                     if peer and persistent else
                                     (", NCCL MIN all-reduce per iteration" if world > 1 else "")),
                   "planner": "one persistent cooperative launch per plan" if persistent
                              else "one cost launch per iteration (CUDA graph)",
                   "l2": "flushed between steps (256 MiB write); within a step the grid is "
                         "L2-resident by design"},
        "iterations_per_sec": iters / (ms_step * 1e-3),
        "e2e": {"value": wps_step * args.steps / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": int(h_start.nbytes + h_init.nbytes),
                "d2h_bytes_per_step": int(2 * h_init.nbytes + 16)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": hbm,
                     "unit": "GB/s", "frac": achieved / hbm,
                     "traffic": (NCU_TRAFFIC.get(args.workload) if kernel_name == "cost_kernel"
                                 else None),
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                     "avg_launch_us": per_launch_s * 1e6, "launches_timed": cost_launches,
                     "alg_bytes_per_launch": alg, "units_per_launch": units_per_launch,
                     "kernel_share_of_step": cost_ms / loop_ms if loop_ms else None},
        "clocks": clocks,
        "best_score": r["best_score"],
    }
    gathers_per_s = links * units_per_launch / per_launch_s
    line["gather_roofline"] = {
        "what": "north_star's L2/HBM gather roofline: collision lookups per second of the cost "
                "kernel against the measured rate of independent random one-sector loads "
                "(doapp_probe_gather) from an 8 MiB (L2) and a 1 GiB (HBM) array",
        "achieved_gathers_per_s": gathers_per_s, "peak_l2_gathers_per_s": g_l2,
        "peak_hbm_gathers_per_s": g_hbm, "regime": "l2" if g_peak is g_l2 else "hbm",
        "frac": gathers_per_s / g_peak,
        "note": "the kernel is issue-bound (about 290 instructions per waypoint-sample for 5 "
                "lookups), not gather-bound: see DESIGN.md section 5"}
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        cpu_iters = iters
        val, secs, sample = cpu_plan_rate(w, cpu_iters, threads, min_seconds=10.0)
        line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": sample, "seconds": secs}
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w, rank)
    else:
        if world != args.gpus and world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun for --gpus > 1")
        if args.workload == "c3":
            run_c3(args, w, rank, world, local_rank)
        else:
            run_ours(args, w, rank, world, local_rank)


if __name__ == "__main__":
    main()
