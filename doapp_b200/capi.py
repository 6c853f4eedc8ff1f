"""ctypes binding of include/doapp/doapp.h.

This is the Python-side stub a test or a bench uses to reach the C ABI; it adds
no logic of its own (numpy arrays in, numpy arrays out) and never falls back to
a CPU path: if the shared library is missing or no CUDA device is present the
calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libdoapp_b200.so")

OK, INVALID_ARG, CUDA_ERROR, NCCL_ERROR, OOM, NOT_READY = range(6)
COLLISION_AS_SHIPPED, COLLISION_INTENDED = 0, 1
SELECT_CEIL_INT, SELECT_FLOAT = 0, 1
UPDATE_KEEP_BEST, UPDATE_WEIGHTED = 0, 1
GRID_NONE, GRID_U8, GRID_F32 = 0, 1, 2
FRAME_CDG, FRAME_OG = 0, 1
COMM_ID_BYTES = 128
PEER_HANDLE_BYTES = 64

# every symbol include/doapp/doapp.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "doapp_config_init", "doapp_version", "doapp_status_str", "doapp_create", "doapp_destroy",
    "doapp_last_error", "doapp_grid_alloc", "doapp_grid_upload", "doapp_grid_download",
    "doapp_grid_from_msg", "doapp_grid_from_depth", "doapp_grid_distance_field", "doapp_distance_field_noalloc", "doapp_grid_scatter", "doapp_grid_scatter_dev", "doapp_grid_lookup",
    "doapp_set_limits", "doapp_fk_model_rexarm", "doapp_set_fk_model", "doapp_noise_generate", "doapp_cost_eval", "doapp_stage_inputs",
    "doapp_cost_eval_dev", "doapp_read_results", "doapp_plan_lerp", "doapp_plan",
    "doapp_plan_async", "doapp_plan_wait", "doapp_plan_state", "doapp_replan_dynamic",
    "doapp_plan_bench", "doapp_plan_profile", "doapp_probe_gather", "doapp_peer_export", "doapp_peer_import", "doapp_comm_unique_id", "doapp_comm_init", "doapp_grid_broadcast",
    "doapp_plan_mode", "doapp_plan_stats", "doapp_launch_count", "doapp_debug_sincos", "doapp_debug_fk",
    "doapp_debug_fk_model",
]


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32), ("num_climbers", C.c_uint32),
        ("samples_per_climber", C.c_uint32), ("num_waypoints", C.c_uint32),
        ("waypoint_dim", C.c_uint32), ("noise_waypoints", C.c_uint32), ("delta_t", C.c_float),
        ("smooth_scale", C.c_float), ("collision_distance", C.c_float),
        ("collision_mode", C.c_int32), ("select_mode", C.c_int32), ("update_mode", C.c_int32),
        ("update_gain", C.c_float), ("limit_weight", C.c_float), ("sigma", C.c_float),
        ("seed", C.c_uint64), ("sample_offset", C.c_uint64),
        ("global_samples_per_climber", C.c_uint64),
    ]


FK_MAX_SPHERES = 16


class FkSphere(C.Structure):
    _fields_ = [("link", C.c_int32), ("centre", C.c_float * 3), ("radius", C.c_float)]


class FkModel(C.Structure):
    """doapp_fk_model: URDF joint origins / axes + per-link collision spheres."""
    _fields_ = [("struct_size", C.c_uint32), ("joint_origin", (C.c_float * 3) * 5),
                ("joint_axis", C.c_int32 * 5), ("num_spheres", C.c_uint32),
                ("spheres", FkSphere * FK_MAX_SPHERES)]


class DoappError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"doapp status {status}: {message}")
        self.status = status
        self.message = message


_lib = None


def load_library(path: str | None = None) -> C.CDLL:
    """dlopen the in-tree CUDA library; raises if it has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("DOAPP_LIB") or LIB_PATH  # DOAPP_LIB: tuning builds only
    if not os.path.exists(p):
        raise FileNotFoundError(
            f"{p} not found: build it with `python -m doapp_b200.build` (no CPU fallback exists)")
    L = C.CDLL(p)
    vp, f32p = C.c_void_p, C.c_void_p
    L.doapp_config_init.argtypes = [C.POINTER(Config)]
    L.doapp_config_init.restype = None
    L.doapp_version.restype = C.c_uint32
    L.doapp_status_str.argtypes = [C.c_int]
    L.doapp_status_str.restype = C.c_char_p
    L.doapp_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.doapp_destroy.argtypes = [vp]
    L.doapp_destroy.restype = None
    L.doapp_last_error.argtypes = [vp]
    L.doapp_last_error.restype = C.c_char_p
    L.doapp_grid_alloc.argtypes = [vp, C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, C.c_uint32,
                                   C.c_double]
    L.doapp_grid_upload.argtypes = [vp, vp, C.c_size_t]
    L.doapp_grid_download.argtypes = [vp, vp, C.c_size_t]
    L.doapp_grid_from_msg.argtypes = [vp, C.c_double, C.c_double, vp, C.c_size_t]
    L.doapp_grid_scatter.argtypes = [vp, f32p, C.c_size_t, C.c_int8, C.c_int]
    L.doapp_grid_scatter_dev.argtypes = [vp, vp, C.c_size_t, C.c_int8, C.c_int]
    L.doapp_grid_lookup.argtypes = [vp, f32p, C.c_size_t, vp, vp]
    L.doapp_set_limits.argtypes = [vp, f32p, f32p]
    L.doapp_fk_model_rexarm.argtypes = [C.POINTER(FkModel)]
    L.doapp_fk_model_rexarm.restype = None
    L.doapp_set_fk_model.argtypes = [vp, C.POINTER(FkModel)]
    L.doapp_debug_fk_model.argtypes = [vp, f32p, C.c_size_t, f32p]
    L.doapp_noise_generate.argtypes = [vp, C.c_uint32, f32p]
    L.doapp_cost_eval.argtypes = [vp, f32p, f32p, f32p, C.c_uint32, f32p, vp, vp, vp]
    L.doapp_stage_inputs.argtypes = [vp, f32p, f32p, f32p]
    L.doapp_cost_eval_dev.argtypes = [vp, C.c_uint32, C.c_int, C.c_uint32, vp]
    L.doapp_read_results.argtypes = [vp, f32p, vp, vp, vp]
    L.doapp_plan_lerp.argtypes = [f32p, f32p, C.c_uint32, C.c_uint32, f32p]
    L.doapp_plan.argtypes = [vp, f32p, f32p, f32p, C.c_uint32, f32p, f32p, f32p, vp, vp]
    L.doapp_plan_async.argtypes = [vp, f32p, f32p, f32p, C.c_uint32, f32p]
    L.doapp_plan_wait.argtypes = [vp, f32p, f32p, vp, vp]
    L.doapp_plan_state.argtypes = [vp, f32p, f32p, vp, vp]
    L.doapp_replan_dynamic.argtypes = [vp, f32p, f32p, f32p, C.c_uint32, f32p, C.c_size_t, f32p,
                                       vp, vp]
    L.doapp_plan_bench.argtypes = [vp, C.c_uint32, C.c_uint32, vp]
    L.doapp_plan_profile.argtypes = [vp, C.c_uint32, vp, vp, vp]
    L.doapp_probe_gather.argtypes = [vp, C.c_size_t, C.c_uint32, vp]
    L.doapp_grid_from_depth.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, C.c_double,
                                        C.c_int8, C.c_int, vp]
    L.doapp_grid_distance_field.argtypes = [vp]
    L.doapp_distance_field_noalloc.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_double,
                                               vp, vp]
    L.doapp_peer_export.argtypes = [vp, vp]
    L.doapp_peer_import.argtypes = [vp, vp, C.c_int32, C.c_int32]
    L.doapp_comm_unique_id.argtypes = [vp]
    L.doapp_comm_init.argtypes = [vp, vp, C.c_int32, C.c_int32]
    L.doapp_grid_broadcast.argtypes = [vp, C.c_int32]
    L.doapp_plan_mode.argtypes = [vp]
    L.doapp_plan_mode.restype = C.c_uint32
    L.doapp_plan_stats.argtypes = [vp, vp, vp]
    L.doapp_launch_count.argtypes = [vp]
    L.doapp_launch_count.restype = C.c_uint64
    L.doapp_debug_sincos.argtypes = [vp, f32p, C.c_size_t, f32p, f32p]
    L.doapp_debug_fk.argtypes = [vp, f32p, C.c_size_t, f32p]
    if path is None:
        _lib = L
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def fk_model_rexarm() -> FkModel:
    m = FkModel()
    load_library().doapp_fk_model_rexarm(C.byref(m))
    return m


def default_config() -> Config:
    cfg = Config()
    load_library().doapp_config_init(C.byref(cfg))
    return cfg


def plan_lerp(start, goal, n: int) -> np.ndarray:
    start, goal = _f32(start), _f32(goal)
    out = np.empty((n, start.shape[0]), np.float32)
    st = load_library().doapp_plan_lerp(_p(start), _p(goal), n, start.shape[0], _p(out))
    if st != OK:
        raise DoappError(st, "doapp_plan_lerp")
    return out


def comm_unique_id() -> bytes:
    buf = (C.c_uint8 * COMM_ID_BYTES)()
    st = load_library().doapp_comm_unique_id(buf)
    if st != OK:
        raise DoappError(st, load_library().doapp_last_error(None).decode())
    return bytes(buf)


class Context:
    """One doapp_ctx: owns the device buffers of one (k, m, n, d) problem on one GPU."""

    def __init__(self, k=1, m=1024, n=64, d=5, n_noise=1, device=0, **kw):
        self.L = load_library()
        cfg = default_config()
        cfg.device = device
        cfg.num_climbers, cfg.samples_per_climber = k, m
        cfg.num_waypoints, cfg.waypoint_dim, cfg.noise_waypoints = n, d, n_noise
        for key, val in kw.items():
            if not hasattr(cfg, key):
                raise AttributeError(key)
            setattr(cfg, key, val)
        self.cfg = cfg
        self.k, self.m, self.n, self.d, self.n_noise = k, m, n, d, n_noise
        self.m_global = cfg.global_samples_per_climber or m
        h = C.c_void_p()
        st = self.L.doapp_create(C.byref(cfg), C.byref(h))
        if st != OK:
            raise DoappError(st, self.L.doapp_last_error(None).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.doapp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, st):
        if st != OK:
            raise DoappError(st, self.L.doapp_last_error(self.h).decode())

    # -- world model -------------------------------------------------------
    def set_grid(self, cells: np.ndarray, resolution: float, frame=FRAME_CDG):
        """cells indexed [iz, iy, ix]; uint8/int8 -> occupancy, float32 -> distance."""
        assert cells.ndim == 3
        kind = GRID_F32 if cells.dtype == np.float32 else GRID_U8
        cells = np.ascontiguousarray(cells if kind == GRID_F32 else cells.astype(np.uint8))
        nz, ny, nx = cells.shape
        self._ck(self.L.doapp_grid_alloc(self.h, kind, frame, nx, ny, nz, resolution))
        self._ck(self.L.doapp_grid_upload(self.h, _p(cells), cells.size))

    def grid_alloc(self, kind, frame, nx, ny, nz, resolution):
        self._ck(self.L.doapp_grid_alloc(self.h, kind, frame, nx, ny, nz, resolution))
        self._grid_shape = (nz, ny, nx)
        self._grid_dtype = np.float32 if kind == GRID_F32 else np.uint8

    def grid_from_msg(self, size: float, granularity: float, grid: np.ndarray):
        grid = np.ascontiguousarray(grid, dtype=np.int8)
        self._ck(self.L.doapp_grid_from_msg(self.h, size, granularity, _p(grid), grid.size))

    def grid_download(self, shape, dtype=np.uint8) -> np.ndarray:
        out = np.empty(shape, dtype)
        self._ck(self.L.doapp_grid_download(self.h, _p(out), out.size))
        return out

    def scatter(self, xyz, val=1, clear_first=True):
        xyz = _f32(xyz).reshape(-1, 3)
        self._ck(self.L.doapp_grid_scatter(self.h, _p(xyz), xyz.shape[0], val, int(clear_first)))

    def grid_lookup(self, xyz):
        xyz = _f32(xyz).reshape(-1, 3)
        idx = np.empty(xyz.shape[0], np.int64)
        col = np.empty(xyz.shape[0], np.uint8)
        self._ck(self.L.doapp_grid_lookup(self.h, _p(xyz), xyz.shape[0], _p(idx), _p(col)))
        return idx, col

    def set_limits(self, lo, hi):
        lo, hi = _f32(lo), _f32(hi)
        self._ck(self.L.doapp_set_limits(self.h, _p(lo), _p(hi)))

    def set_fk_model(self, model):
        """FkModel -> URDF-correct chain with per-link spheres (rexarm.urdf:61-149);
        None -> the chain of generate_trajectories.cu:397-463."""
        self._fk_spheres = model.num_spheres if model is not None else 0
        self._ck(self.L.doapp_set_fk_model(self.h, C.byref(model) if model is not None else None))

    def debug_fk_model(self, waypoints):
        wp = _f32(waypoints).reshape(-1, self.d)
        out = np.empty((wp.shape[0], self._fk_spheres, 3), np.float32)
        self._ck(self.L.doapp_debug_fk_model(self.h, _p(wp), wp.shape[0], _p(out)))
        return out

    # -- stages ------------------------------------------------------------
    def noise(self, iteration=0) -> np.ndarray:
        out = np.empty((self.k, self.m, self.n_noise, self.d), np.float32)
        self._ck(self.L.doapp_noise_generate(self.h, iteration, _p(out)))
        return out

    def cost_eval(self, base_traj, start, noise=None, iteration=0):
        base_traj, start, noise = _f32(base_traj), _f32(start), _f32(noise)
        assert base_traj.size == self.k * self.n * self.d
        if noise is not None:
            assert noise.size == self.k * self.m * self.n_noise * self.d
        scores = np.empty((self.k, self.m), np.float32)
        coll = np.empty((self.k, self.m), np.int32)
        am, mn = C.c_int64(), C.c_float()
        self._ck(self.L.doapp_cost_eval(self.h, _p(base_traj), _p(start), _p(noise), iteration,
                                        _p(scores), _p(coll), C.addressof(am), C.addressof(mn)))
        return dict(scores=scores, collisions=coll, argmin=am.value, min_score=mn.value)

    def stage_inputs(self, base_traj, start, noise=None):
        base_traj, start, noise = _f32(base_traj), _f32(start), _f32(noise)
        self._ck(self.L.doapp_stage_inputs(self.h, _p(base_traj), _p(start), _p(noise)))

    def cost_eval_dev(self, iteration=0, use_staged_noise=False, repeats=1) -> float:
        ms = C.c_float()
        self._ck(self.L.doapp_cost_eval_dev(self.h, iteration, int(use_staged_noise), repeats,
                                            C.addressof(ms)))
        return ms.value

    def read_results(self, want_arrays=True):
        scores = np.empty((self.k, self.m), np.float32) if want_arrays else None
        coll = np.empty((self.k, self.m), np.int32) if want_arrays else None
        am, mn = C.c_int64(), C.c_float()
        self._ck(self.L.doapp_read_results(self.h, _p(scores), _p(coll), C.addressof(am),
                                           C.addressof(mn)))
        return dict(scores=scores, collisions=coll, argmin=am.value, min_score=mn.value)

    # -- planner -----------------------------------------------------------
    def plan(self, start, goal, init_traj, iters, noise=None):
        self.plan_async(start, goal, init_traj, iters, noise)
        return self.plan_wait()

    def plan_async(self, start, goal, init_traj, iters, noise=None):
        start, goal, init_traj, noise = _f32(start), _f32(goal), _f32(init_traj), _f32(noise)
        assert init_traj.size == self.n * self.d
        if noise is not None:
            assert noise.size == iters * self.k * self.m_global * self.n_noise * self.d
        self._keep = (start, goal, init_traj, noise)
        self._ck(self.L.doapp_plan_async(self.h, _p(start), _p(goal), _p(init_traj), iters,
                                         _p(noise)))

    def plan_wait(self):
        bt, bv = np.empty((self.n, self.d), np.float32), np.empty((self.n, self.d), np.float32)
        bs, bc = C.c_float(), C.c_int64()
        self._ck(self.L.doapp_plan_wait(self.h, _p(bt), _p(bv), C.addressof(bs), C.addressof(bc)))
        return dict(best_traj=bt, best_vel=bv, best_score=bs.value, best_climber=bc.value)

    def plan_state(self):
        tr = np.empty((self.k, self.n, self.d), np.float32)
        sc, rd = np.empty(self.k, np.float32), np.empty(self.k, np.int32)
        fb = np.empty(self.k, np.uint8)
        self._ck(self.L.doapp_plan_state(self.h, _p(tr), _p(sc), _p(rd), _p(fb)))
        return dict(trajectories=tr, scores=sc, rounded=rd, found_better=fb)

    def plan_bench(self, iters, repeats=1) -> float:
        ms = C.c_float()
        self._ck(self.L.doapp_plan_bench(self.h, iters, repeats, C.addressof(ms)))
        return ms.value

    def plan_profile(self, iters):
        """(summed cost-kernel ms, launches, loop ms) of one event-instrumented pass."""
        tot, lp, cnt = C.c_float(), C.c_float(), C.c_uint32()
        self._ck(self.L.doapp_plan_profile(self.h, iters, C.addressof(tot), C.addressof(cnt),
                                           C.addressof(lp)))
        return tot.value, cnt.value, lp.value

    def replan_dynamic(self, start, goal, init_traj, clouds, staged_shape=None):
        """clouds [iters][P][3] host array; or None with staged_shape = (iters, P): reuse the
        clouds the previous call left in device memory (no host-to-device copy)."""
        start, goal, init_traj = _f32(start), _f32(goal), _f32(init_traj)
        if clouds is None:
            iters, npts = staged_shape
        else:
            clouds = _f32(clouds)
            iters, npts = clouds.shape[0], clouds.shape[1]
        bt = np.empty((self.n, self.d), np.float32)
        bs, ms = C.c_float(), C.c_float()
        self._ck(self.L.doapp_replan_dynamic(self.h, _p(start), _p(goal), _p(init_traj), iters,
                                             _p(clouds), npts, _p(bt), C.addressof(bs),
                                             C.addressof(ms)))
        return dict(best_traj=bt, best_score=bs.value, elapsed_ms=ms.value)

    def grid_from_depth(self, depth, intrinsics, extrinsics, crop_size, val=1, clear_first=True) -> int:
        """depth [cams][h][w] uint16 mm; intrinsics [cams][4] f64 fx,fy,cx,cy; extrinsics [cams][12] f32.
        Returns the number of points that passed the crop box."""
        depth = np.ascontiguousarray(depth, np.uint16)
        intr = np.ascontiguousarray(intrinsics, np.float64)
        extr = np.ascontiguousarray(extrinsics, np.float32)
        cams, h, w = depth.shape
        kept = C.c_uint64()
        self._ck(self.L.doapp_grid_from_depth(self.h, _p(depth), cams, w, h, _p(intr), _p(extr),
                                              float(crop_size), val, int(clear_first), C.addressof(kept)))
        return int(kept.value)

    def grid_distance_field(self):
        """U8 occupancy -> F32 exact Euclidean distance field (same frame and dimensions)."""
        self._ck(self.L.doapp_grid_distance_field(self.h))

    @staticmethod
    def distance_field_noalloc(occ_dev_ptr: int, nx: int, ny: int, nz: int, resolution: float,
                               dist_dev_ptr: int, stream: int = 0):
        """doapp_distance_field_noalloc: device pointers in and out, static scratch inside the
        library, no allocation; enqueues on `stream` (a cudaStream_t value, 0 = default)."""
        L = load_library()
        st = L.doapp_distance_field_noalloc(C.c_void_p(occ_dev_ptr), nx, ny, nz, float(resolution),
                                            C.c_void_p(dist_dev_ptr), C.c_void_p(stream or None))
        if st != OK:
            raise DoappError(st, L.doapp_last_error(None).decode())

    def probe_gather(self, array_bytes: int, cell_bytes: int = 1) -> float:
        """Random single-cell loads per second from a scratch array (roofline denominator)."""
        out = C.c_double()
        self._ck(self.L.doapp_probe_gather(self.h, array_bytes, cell_bytes, C.addressof(out)))
        return out.value

    # -- multi-GPU ---------------------------------------------------------
    def comm_init(self, uid: bytes, rank: int, nranks: int):
        buf = (C.c_uint8 * COMM_ID_BYTES).from_buffer_copy(uid)
        self._ck(self.L.doapp_comm_init(self.h, buf, rank, nranks))

    def peer_export(self) -> bytes:
        buf = (C.c_uint8 * PEER_HANDLE_BYTES)()
        self._ck(self.L.doapp_peer_export(self.h, buf))
        return bytes(buf)

    def peer_import(self, handles: bytes, rank: int, nranks: int):
        assert len(handles) == PEER_HANDLE_BYTES * nranks
        buf = (C.c_uint8 * len(handles)).from_buffer_copy(handles)
        self._ck(self.L.doapp_peer_import(self.h, buf, rank, nranks))

    def peer_disable(self):
        """Back to the NCCL per-iteration merge (a peer failed to map the mailboxes)."""
        self._ck(self.L.doapp_peer_import(self.h, None, 0, 0))

    def grid_broadcast(self, root=0):
        self._ck(self.L.doapp_grid_broadcast(self.h, root))

    # -- introspection -----------------------------------------------------
    def plan_mode(self) -> int:
        """bit 0: persistent one-launch planner; bit 1: in-kernel peer-mailbox argmin merge."""
        return int(self.L.doapp_plan_mode(self.h))

    def plan_stats(self) -> dict:
        acc, passes = C.c_uint64(), C.c_uint64()
        self._ck(self.L.doapp_plan_stats(self.h, C.addressof(acc), C.addressof(passes)))
        return dict(accepted_updates=acc.value, cost_passes=passes.value)

    def launch_count(self) -> int:
        return int(self.L.doapp_launch_count(self.h))

    def debug_sincos(self, x):
        x = _f32(np.atleast_1d(x))
        s, c = np.empty_like(x), np.empty_like(x)
        self._ck(self.L.doapp_debug_sincos(self.h, _p(x), x.size, _p(s), _p(c)))
        return s, c

    def debug_fk(self, waypoints):
        wp = _f32(waypoints).reshape(-1, self.d)
        out = np.empty((wp.shape[0], 5, 3), np.float32)
        self._ck(self.L.doapp_debug_fk(self.h, _p(wp), wp.shape[0], _p(out)))
        return out
