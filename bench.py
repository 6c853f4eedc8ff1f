#!/usr/bin/env python
"""bench.py — throughput of the perturbator hot path on B200.

Default workload (`c4s`): BASELINE config 4 at its largest size, STRONG-scaled — 1 048 576
perturbation samples x 128 waypoints in total (Rexarm 5-DOF, 128^3 uint8 grid, 1 cm), the samples
split evenly over the N GPUs, 20 optimiser iterations per step (noise -> FK -> collision gathers
-> smoothness -> argmin -> keep-best update, one cross-GPU argmin merge per iteration).
`value` = waypoint-samples costed per second over all ranks, inputs resident in HBM, timed with
CUDA events on the ctx stream (max over ranks); `e2e` = the same through the host-buffer C-ABI
call doapp_plan (H2D of start / initial trajectory, D2H of the best trajectory inside the timed
region).  The same JSON line carries, under "also", the other BASELINE configs: `c2` (latency
bound: planner iterations/s, 1024 samples per GPU), and at N = 1 `c5` (512^3 grid) and `c3`
(grid rebuilt from a 300k-point cloud every iteration).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c4s|c2|...]
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
FACTS_FILE = os.path.join(ROOT, "profiles", "r02", "ncu_facts.json")

# BASELINE configs.  `m` = samples per GPU (weak) unless `m_total` is given (strong: split over
# the ranks).
WORKLOADS = {
    "c4s": dict(desc="C4-1M strong-scaled: Rexarm 5-DOF, 128^3 u8 grid (1 cm), 128 waypoints x "
                     "1048576 samples in total split over the GPUs, allreduce-argmin per "
                     "iteration, 20 optimizer iterations per step",
                grid_n=128, res=0.01, n=128, m_total=1048576, iters=20, sigma=0.05, fill=0.0),
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


def shard(w, rank, world):
    """(samples of this rank, its global offset, samples of the whole job)."""
    if "m_total" in w:
        mt = w["m_total"]
        base, extra = divmod(mt, world)
        cnt = base + (1 if rank < extra else 0)
        return cnt, rank * base + min(rank, extra), mt
    return w["m"], rank * w["m"], w["m"] * world


def config_of(w, world):
    """The workload as both arms (ours / reference) name it: identical keys and values."""
    _, _, mt = shard(w, 0, world)
    return {"workload": w["desc"], "samples_total": mt, "waypoints": w["n"], "joints": D,
            "iterations_per_step": w["iters"], "grid": f"{w['grid_n']}^3 u8",
            "lookups_per_waypoint": FK_MODEL_SPHERES if w.get("fk_model") else LINKS,
            "l2": "flushed between steps (256 MiB write); within a step the grid is L2-resident "
                  "by design"}


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


_GRIDS = {}


def make_inputs(w):
    key = (w["grid_n"], w["res"], w["fill"])
    if key not in _GRIDS:
        _GRIDS[key] = synth.box_grid(w["grid_n"], w["res"], fill=w["fill"])
    return _GRIDS[key], synth.START, synth.GOALS[1]


# ------------------------------------------------------------------ CPU legs (the oracle port)
CPU_BUDGET_WPS = 8_388_608  # waypoint-samples per timed CPU repetition (bounded sample)


def cpu_sample_shape(w, world):
    """A bounded sample of the workload for the CPU legs: the first `m` samples of the job for
    `iters` iterations, about CPU_BUDGET_WPS waypoint-samples."""
    _, _, mt = shard(w, 0, world)
    n, iters = w["n"], w["iters"]
    if mt * n * iters > CPU_BUDGET_WPS:
        iters = max(1, min(iters, CPU_BUDGET_WPS // (mt * n)))
    m = max(1, min(mt, CPU_BUDGET_WPS // (n * iters)))
    return m, iters


def cpu_plan_rate(w, m, iters, threads, reps=1, min_seconds=0.0):
    """The oracle's planner loop (CPU restatement of the reference arithmetic) over `m` samples
    x `iters` iterations of workload `w` on `threads` host threads; repeated until `min_seconds`
    of wall time have been spent; returns (wps/s of the best repetition, its seconds, sample)."""
    import oracle
    oracle.build()
    grid, start, goal = make_inputs(w)
    orc = oracle.Oracle(k=1, m=m, n=w["n"], d=D, sigma=w["sigma"], select_mode=1,
                        fk_model=oracle.fk_model_rexarm() if w.get("fk_model") else None)
    orc.set_grid(grid, w["res"])
    init = oracle.plan_lerp(start, goal, w["n"])
    best, total, done = float("inf"), 0.0, 0
    while done < reps or total < min_seconds:
        t0 = time.perf_counter()
        orc.plan(start, init, iters, None, threads=threads)
        dt = time.perf_counter() - t0
        best, total, done = min(best, dt), total + dt, done + 1
    wps = m * w["n"] * iters
    return wps / best, best, (f"first {m} samples x {w['n']} waypoints x {iters} iterations of the "
                              f"workload, best of {done} repetitions ({total:.1f} s)")


def cpu_baseline_of(w, world=1):
    """cpu_baseline object: all host threads and one thread."""
    threads = os.cpu_count() or 1
    m, iters = cpu_sample_shape(w, world)
    val, secs, sample = cpu_plan_rate(w, m, iters, threads, min_seconds=8.0)
    m1 = max(1, m // 8)
    val1, secs1, sample1 = cpu_plan_rate(w, m1, iters, 1, min_seconds=4.0)
    return {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
            "seconds": secs,
            "single_thread": {"value": val1, "unit": UNIT, "cores": 1, "sample": sample1,
                              "seconds": secs1}}


def reference_gpu_kernel():
    """Secondary data point: the reference's OWN kernel (generate_trajectories.cu:144-258,
    compiled unmodified into oracle/_ref/ref_harness) timed on this GPU at the only shape it can
    launch (k=1000 blocks x 1000 threads: m=5, n=50, d=4; collision term stubbed)."""
    import subprocess
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/ref_harness not built"}
    try:
        out = subprocess.run([exe, "--time", "20"], capture_output=True, text=True, timeout=120)
        for ln in out.stdout.splitlines():
            if ln.startswith("{"):
                return json.loads(ln)
        return {"unavailable": (out.stderr or out.stdout)[-200:]}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"}


def run_reference(args, w, rank, world):
    """--impl reference: the reference's CPU cost path.  The reference has no runnable CPU (or
    GPU) implementation of the full path (collision term stubbed), so this is the
    oracle port, threaded over samples the way cpu_distance_grid.cu:88-110 threads its fill.
    Each step costs a bounded sample of the arm's workload (cpu_sample_shape)."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    m, iters = cpu_sample_shape(w, world)
    for _ in range(args.warmup):
        cpu_plan_rate(w, max(1, m // 8), 1, threads)
    secs = []
    for _ in range(args.steps):
        _, s, sample = cpu_plan_rate(w, m, iters, threads)
        secs.append(s)
    ms = 1e3 * float(np.mean(secs))
    val = m * w["n"] * iters / (ms * 1e-3)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "strong" if "m_total" in w else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(w, world),
            "note": "CPU oracle port on the host cores; every step is a bounded sample of the "
                    "workload (throughput, not job time)",
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": sample + " per step"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if world == 1 and not args.no_cpu:
        try:
            import torch
            if torch.cuda.is_available():
                line["reference_gpu_kernel"] = reference_gpu_kernel()
        except Exception:  # noqa: BLE001
            pass
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ our arm
class Dist:
    """torch.distributed plumbing (rendezvous, barrier, max over ranks); no data-path use."""

    def __init__(self, rank, world, local_rank):
        import torch
        self.torch, self.rank, self.world = torch, rank, world
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        self.dist = None
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2

    def barrier(self):
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max(self, x: float) -> float:
        if not self.dist:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def flush_l2(self):
        self.flush.fill_(1)
        self.torch.cuda.synchronize()

    def same_everywhere(self, arr) -> bool:
        if not self.dist:
            return True
        t = self.torch.from_numpy(np.ascontiguousarray(arr).view(np.int32).copy()).to(self.dev)
        lo, hi = t.clone(), t.clone()
        self.dist.all_reduce(lo, op=self.dist.ReduceOp.MIN)
        self.dist.all_reduce(hi, op=self.dist.ReduceOp.MAX)
        return bool(self.torch.equal(lo, hi))

    def close(self):
        if self.dist:
            self.dist.destroy_process_group()


def load_facts():
    try:
        return json.load(open(FACTS_FILE))
    except Exception:  # noqa: BLE001
        return {}


def measure_plan(key, w, dd: Dist, local_rank, steps, warmup, full):
    """One workload through the planner: device-timed value, e2e through doapp_plan, the dominant
    kernel's rooflines, and (N > 1) parity of the sharded result with the 1-GPU answer."""
    import torch
    from doapp_b200 import capi
    rank, world, dev = dd.rank, dd.world, dd.dev
    n, iters = w["n"], w["iters"]
    m, offset, m_total = shard(w, rank, world)
    grid, start, goal = make_inputs(w)
    init = capi.plan_lerp(start, goal, n)
    ctx = capi.Context(k=1, m=m, n=n, d=D, device=local_rank, sigma=w["sigma"], select_mode=1,
                       sample_offset=offset, global_samples_per_climber=m_total)
    if world > 1:
        uid = torch.zeros(capi.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            uid = torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8).to(dev)
        dd.dist.broadcast(uid, 0)
        ctx.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
        # the grid is broadcast once over NVLink (north_star); rank 0 holds the real cells
        if rank == 0:
            ctx.set_grid(grid, w["res"])
        else:
            ctx.grid_alloc(capi.GRID_U8, capi.FRAME_CDG, grid.shape[2], grid.shape[1],
                           grid.shape[0], w["res"])
        ctx.grid_broadcast(0)
        from doapp_b200 import sharding
        if not os.environ.get("DOAPP_NO_PEER"):
            sharding.connect_peers(ctx, rank, world, dev)
    else:
        ctx.set_grid(grid, w["res"])
    links = LINKS
    if w.get("fk_model"):
        ctx.set_fk_model(capi.fk_model_rexarm())
        links = FK_MODEL_SPHERES

    # pinned host staging for the e2e leg
    h_start = torch.from_numpy(start.copy()).pin_memory().numpy()
    h_goal = torch.from_numpy(goal.copy()).pin_memory().numpy()
    h_init = torch.from_numpy(init.copy()).pin_memory().numpy()

    first = ctx.plan(h_start, h_goal, h_init, iters)  # stages inputs, builds the CUDA graph
    first_state = ctx.plan_state()
    for _ in range(warmup):
        ctx.plan_bench(iters, 1)

    # ---- device-timed region: K steps, inputs resident in HBM -------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    dd.barrier()
    l0 = ctx.launch_count()
    ms_total = 0.0
    for _ in range(steps):
        dd.flush_l2()
        ms_total += ctx.plan_bench(iters, 1)
    launches = ctx.launch_count() - l0
    persistent = bool(ctx.plan_mode() & 1)
    peer = bool(ctx.plan_mode() & 2)
    dd.barrier()
    ms_total = dd.max(ms_total)
    clocks = sampler.stop()
    res = ctx.plan_wait()
    assert np.array_equal(res["best_traj"], first["best_traj"]), "bench result not reproducible"
    stats = ctx.plan_stats() if persistent else None

    # ---- e2e: host buffers through doapp_plan, copies inside the timed region ---------
    for _ in range(warmup):
        ctx.plan(h_start, h_goal, h_init, iters)
    dd.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        r = ctx.plan(h_start, h_goal, h_init, iters)
    torch.cuda.synchronize()
    e2e_s = dd.max(time.perf_counter() - t0)
    dd.barrier()

    # ---- dominant kernel: per-launch CUDA events around every cost launch -------------
    cost_ms, cost_launches, loop_ms = ctx.plan_profile(iters)
    cost_ms, cost_launches, loop_ms = ctx.plan_profile(iters)
    ctx.plan_wait()

    wps_step = m_total * n * iters
    ms_step = ms_total / steps
    out = {
        "value": wps_step / (ms_step * 1e-3), "unit": UNIT, "ms_per_step": ms_step,
        "iterations_per_sec": iters / (ms_step * 1e-3),
        "us_per_iteration": 1e3 * ms_step / iters,
        "config": config_of(w, world),
        "run": {"samples_this_gpu": m,
                "parallelism": f"samples sharded over {world} GPU(s), grid replicated"
                               + (", argmin keys merged in-kernel through NVLink peer mailboxes"
                                  if peer else
                                  (", NCCL MIN all-reduce per iteration" if world > 1 else "")),
                "planner": "one persistent cooperative launch per plan" if persistent
                           else "one cost launch per iteration (CUDA graph)"},
        "e2e": {"value": wps_step * steps / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": int(h_start.nbytes + h_init.nbytes),
                "d2h_bytes_per_step": int(2 * h_init.nbytes + 16)},
        "gpu_launches": int(launches),
        "clocks": clocks, "best_score": r["best_score"],
    }
    if stats and stats["cost_passes"]:
        # the persistent planner costs two iterations per pass, speculatively: passes beyond
        # iterations / 2 were voided by accepted keep-best updates (mis-speculation)
        out["planner_stats"] = dict(stats, iterations=iters,
                                    passes_per_iteration=stats["cost_passes"] / iters)

    # ---- rooflines of the dominant kernel ---------------------------------------------
    peaks = {}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    per_launch_s = cost_ms * 1e-3 / max(1, cost_launches)
    # one launch per iteration, or ONE persistent launch running all the iterations
    units_per_launch = m * n * iters / max(1, cost_launches)
    alg = alg_bytes_per_wps(n, links=links) * units_per_launch
    kernel_name = "plan_persistent_kernel" if persistent else "cost_kernel_pk / cost_kernel"
    if w.get("fk_model"):
        kernel_name += "<FK model>"
    achieved = alg / per_launch_s / 1e9
    fact = load_facts().get(key) or {}
    traffic = None
    if fact.get("dram_bytes_per_launch") is not None and fact.get("units_per_launch"):
        traffic = fact["dram_bytes_per_launch"] * units_per_launch / fact["units_per_launch"]
    out["roofline"] = {
        "bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": hbm, "unit": "GB/s",
        "frac": achieved / hbm, "traffic": traffic,
        "traffic_source": fact.get("source"),
        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
        "avg_launch_us": per_launch_s * 1e6, "launches_timed": cost_launches,
        "alg_bytes_per_launch": alg, "units_per_launch": units_per_launch,
        "kernel_share_of_step": cost_ms / loop_ms if loop_ms else None,
        "note": "DRAM traffic is far below the algorithmic bytes: the grid and the trajectory are "
                "L1/L2-resident, the kernel is bound by instruction issue (see `issue`)"}
    if fact.get("warp_inst_per_wps"):
        sm_mhz = clocks.get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        peak_issue = sms * 4 * sm_mhz * 1e6                     # warp instructions / s
        ach_issue = fact["warp_inst_per_wps"] * units_per_launch / per_launch_s
        out["roofline"]["issue"] = {
            "what": "the binding roofline: warp instructions issued per second by the dominant "
                    "kernel (instruction count per waypoint-sample from the committed ncu "
                    "capture x measured rate) against SMs x 4 schedulers x SM clock",
            "thread_inst_per_wps": fact["thread_inst_per_wps"],
            "achieved_warp_inst_per_s": ach_issue, "peak_warp_inst_per_s": peak_issue,
            "frac": ach_issue / peak_issue, "sm_mhz": sm_mhz, "sms": sms,
            "wps_per_s_at_full_issue": peak_issue / fact["warp_inst_per_wps"],
            "inst_count_source": fact.get("source")}
    if full:
        # measured rate of independent random one-sector loads, L2- and HBM-resident
        g_l2 = ctx.probe_gather(8 << 20, 1)
        g_hbm = ctx.probe_gather(1 << 30, 1)
        gathers_per_s = links * units_per_launch / per_launch_s
        out["gather_rate"] = {
            "what": "collision lookups per second of the cost kernel beside the measured rate of "
                    "independent RANDOM one-sector loads (doapp_probe_gather, 8 MiB = L2, 1 GiB "
                    "= HBM).  Not a bound for this kernel: neighbouring samples look up "
                    "neighbouring voxels (coherent lookups hit L1), so the ratio may exceed 1",
            "achieved_lookups_per_s": gathers_per_s, "random_l2_gathers_per_s": g_l2,
            "random_hbm_gathers_per_s": g_hbm,
            "coherent_vs_random_l2": gathers_per_s / g_l2,
            "coherent_vs_random_hbm": gathers_per_s / g_hbm}

    # ---- N > 1: the sharded plan must equal the 1-GPU plan over the global sample set ---
    if world > 1:
        same = dd.same_everywhere(first_state["trajectories"]) and dd.same_everywhere(
            np.array([first["best_score"]], np.float32))
        ok = True
        if rank == 0:
            one = capi.Context(k=1, m=m_total, n=n, d=D, device=local_rank, sigma=w["sigma"],
                               select_mode=1)
            one.set_grid(grid, w["res"])
            if w.get("fk_model"):
                one.set_fk_model(capi.fk_model_rexarm())
            r1 = one.plan(h_start, h_goal, h_init, iters)
            s1 = one.plan_state()
            one.close()
            ok = (np.array_equal(r1["best_traj"].view(np.uint32), first["best_traj"].view(np.uint32))
                  and np.float32(r1["best_score"]).view(np.uint32) == np.float32(first["best_score"]).view(np.uint32)
                  and np.array_equal(s1["trajectories"].view(np.uint32),
                                     first_state["trajectories"].view(np.uint32)))
        out["parity_vs_1gpu"] = bool(same and ok)
        out["ranks_agree"] = bool(same)
        assert same and ok, f"{key}: sharded plan differs from the 1-GPU plan (ranks_agree={same})"
    ctx.close()
    return out


def measure_c3(w, dd: Dist, local_rank, steps, warmup):
    """BASELINE config 3: every iteration rebuilds the grid from a fresh cloud (stage 5:
    cudaMemset clear + scatter_kernel), re-scores the kept trajectory and runs one optimizer
    iteration (doapp_replan_dynamic).  Replicas only at N > 1 (every rank rebuilds its own copy
    of the grid from the same cloud: cheaper than broadcasting it)."""
    import torch
    from doapp_b200 import capi
    m, n, iters, P = w["m"], w["n"], w["iters"], w["points"]
    grid, start, goal = make_inputs(w)
    init = capi.plan_lerp(start, goal, n)
    ctx = capi.Context(k=1, m=m, n=n, d=D, device=local_rank, sigma=w["sigma"], select_mode=1)
    ctx.set_grid(np.zeros_like(grid), w["res"])
    clouds = np.stack([synth.cloud(P, it, size=w["grid_n"] * w["res"]) for it in range(iters)])
    h_clouds = torch.from_numpy(clouds).pin_memory().numpy()
    for _ in range(max(3, warmup)):
        ctx.replan_dynamic(start, goal, init, h_clouds)
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = ctx.launch_count()
    dev_ms = 0.0
    for _ in range(steps):  # clouds resident in HBM (staged by the warm-up calls): CUDA events
        r = ctx.replan_dynamic(start, goal, init, None, staged_shape=(iters, P))
        dev_ms += r["elapsed_ms"]
    launches = ctx.launch_count() - l0
    t0 = time.perf_counter()
    for _ in range(steps):  # e2e: the clouds streamed in (per-cloud copies overlap the loop) + D2H
        r = ctx.replan_dynamic(start, goal, init, h_clouds)
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()
    wps_step = m * n * iters
    ms_step = dev_ms / steps
    cells = w["grid_n"] ** 3
    alg = 12.0 * P + P + cells          # cloud read + byte stores + clear, per rebuild
    out = {
        "value": wps_step / (ms_step * 1e-3), "unit": UNIT, "ms_per_step": ms_step,
        "config": dict(config_of(w, 1), points_per_cloud=P,
                       l2="cloud of every iteration is new data (72 MB per step streamed from "
                          "HBM; in the e2e leg from pinned host memory, copy of cloud i+1 "
                          "overlapping the replan of cloud i); grid L2-resident by design"),
        "iterations_per_sec": iters / (ms_step * 1e-3),
        "grid_rebuilds_per_sec": iters / (ms_step * 1e-3),
        "points_per_sec": P * iters / (ms_step * 1e-3),
        "e2e": {"value": wps_step * steps / e2e_s, "unit": UNIT,
                "h2d_bytes_per_step": int(h_clouds.nbytes + init.nbytes + start.nbytes),
                "d2h_bytes_per_step": int(init.nbytes + 8)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "scatter_kernel + clear", "achieved": None,
                     "peak": None, "unit": "GB/s", "frac": None, "traffic": None,
                     "alg_bytes_per_launch": alg,
                     "note": "latency-bound: 5.7 MB per rebuild is < 1 us of HBM time"},
        "clocks": clocks, "best_score": r["best_score"],
    }
    ctx.close()
    return out


def brief(o):
    """Sub-object of the `also` block: the numbers, without the prose."""
    keep = ("value", "unit", "ms_per_step", "iterations_per_sec", "us_per_iteration", "e2e",
            "gpu_launches", "parity_vs_1gpu", "grid_rebuilds_per_sec", "points_per_sec",
            "best_score", "planner_stats")
    b = {k: o[k] for k in keep if k in o}
    b["workload"] = o["config"]["workload"]
    if "run" in o:
        b["planner"] = o["run"]["planner"]
        b["parallelism"] = o["run"]["parallelism"]
    rf = o.get("roofline") or {}
    b["roofline"] = {k: rf.get(k) for k in ("kernel", "achieved", "peak", "frac", "traffic",
                                            "avg_launch_us") if k in rf}
    if "issue" in rf:
        b["roofline"]["issue_frac"] = rf["issue"]["frac"]
        b["roofline"]["thread_inst_per_wps"] = rf["issue"]["thread_inst_per_wps"]
    return b


def run_ours(args, rank, world, local_rank):
    dd = Dist(rank, world, local_rank)
    key = args.workload
    w = WORKLOADS[key]
    if key == "c3":
        o = measure_c3(w, dd, local_rank, args.steps, args.warmup)
    else:
        o = measure_plan(key, w, dd, local_rank, args.steps, args.warmup, full=True)
    line = {"metric": METRIC, "value": o["value"], "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": o["ms_per_step"],
            "higher_is_better": True, "scaling": "strong" if "m_total" in w else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    line.update({k: v for k, v in o.items() if k not in line})
    if not args.no_also and key == "c4s":
        also = {}
        sub_steps = max(3, min(args.steps, 10))
        try:
            also["c2"] = brief(measure_plan("c2", WORKLOADS["c2"], dd, local_rank, sub_steps,
                                            max(3, args.warmup), full=False))
            if world == 1:
                also["c5"] = brief(measure_plan("c5", WORKLOADS["c5"], dd, local_rank, sub_steps,
                                                max(3, args.warmup), full=False))
                also["c3"] = brief(measure_c3(WORKLOADS["c3"], dd, local_rank, sub_steps,
                                              args.warmup))
        except AssertionError:
            raise
        except Exception as e:  # noqa: BLE001 - a failing side workload must not lose the headline
            also["error"] = f"{type(e).__name__}: {e}"
        line["also"] = also
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline_of(w, world)
    if rank == 0:
        print(json.dumps(line), flush=True)
    dd.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4s", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-also", action="store_true", help="skip the side workloads (c2, c5, c3)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, WORKLOADS[args.workload], rank, max(world, args.gpus))
    else:
        if world != args.gpus and world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun for --gpus > 1")
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
