"""ctypes binding of the CPU oracle (oracle/doapp_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs, never by doapp_b200.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libdoapp_oracle.so")

COLLISION_AS_SHIPPED, COLLISION_INTENDED = 0, 1
SELECT_CEIL_INT, SELECT_FLOAT = 0, 1
GRID_NONE, GRID_U8, GRID_F32 = 0, 1, 2
FRAME_CDG, FRAME_OG = 0, 1
SUM_SEQUENTIAL, SUM_TREE32, SUM_SEG4, SUM_WAYPOINT = 0, 1, 2, 3


class Grid(C.Structure):
    _fields_ = [("kind", C.c_int32), ("frame", C.c_int32), ("nx", C.c_uint32), ("ny", C.c_uint32),
                ("nz", C.c_uint32), ("resolution", C.c_double), ("cells", C.c_void_p)]


class Params(C.Structure):
    _fields_ = [("k", C.c_uint32), ("m", C.c_uint32), ("n", C.c_uint32), ("d", C.c_uint32),
                ("n_noise", C.c_uint32), ("delta_t", C.c_float), ("smooth_scale", C.c_float),
                ("collision_distance", C.c_float), ("collision_mode", C.c_int32),
                ("select_mode", C.c_int32), ("sum_order", C.c_int32), ("limit_weight", C.c_float),
                ("limit_min", C.c_void_p), ("limit_max", C.c_void_p), ("sigma", C.c_float),
                ("seed", C.c_uint64), ("sample_offset", C.c_uint64), ("m_global", C.c_uint64),
                ("update_mode", C.c_int32), ("update_gain", C.c_float),
                ("fk_model", C.c_void_p)]


FK_MAX_SPHERES = 16


class FkSphere(C.Structure):
    _fields_ = [("link", C.c_int32), ("centre", C.c_float * 3), ("radius", C.c_float)]


class FkModel(C.Structure):
    """orc_fk_model: URDF joint origins / axes + per-link collision spheres."""
    _fields_ = [("joint_origin", (C.c_float * 3) * 5), ("joint_axis", C.c_int32 * 5),
                ("num_spheres", C.c_uint32), ("spheres", FkSphere * FK_MAX_SPHERES)]


def fk_model_rexarm() -> FkModel:
    m = FkModel()
    lib().orc_fk_model_rexarm(C.byref(m))
    return m


def fk_model_points(model: FkModel, waypoints) -> np.ndarray:
    """[count][num_spheres][3] sphere centres of the URDF-correct chain."""
    wp = _f32(waypoints)
    wp = wp.reshape(-1, wp.shape[-1])
    out = np.empty((wp.shape[0], model.num_spheres, 3), np.float32)
    for i in range(wp.shape[0]):
        lib().orc_fk_model_points(C.byref(model), wp[i].ctypes.data_as(C.c_void_p), wp.shape[1],
                                  out[i].ctypes.data_as(C.c_void_p))
    return out


def _digest() -> str:
    import hashlib
    h = hashlib.sha256()
    for f in ("doapp_oracle.c", "doapp_oracle.h", "Makefile"):
        h.update(open(os.path.join(_HERE, f), "rb").read())
    return h.hexdigest()


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (oracle/Makefile).  Staleness is a content hash: file
    mtimes do not survive the copy to the GPU box."""
    stamp = _LIB_PATH + ".stamp"
    fresh = (os.path.exists(_LIB_PATH) and os.path.exists(stamp)
             and open(stamp).read().strip() == _digest())
    if force or not fresh:
        subprocess.run(["make", "-C", _HERE, "-B", "-s", "all"], check=True,
                       stdout=subprocess.DEVNULL)
        with open(stamp, "w") as f:
            f.write(_digest())
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        fp = C.POINTER(C.c_float)
        L.orc_params_init.argtypes = [C.POINTER(Params)]
        L.orc_sincosf.argtypes = [C.c_float, fp, fp]
        L.orc_logf.argtypes = [C.c_float]
        L.orc_logf.restype = C.c_float
        L.orc_philox4x32_10.argtypes = [C.c_void_p] * 3
        L.orc_noise_fill.argtypes = [C.POINTER(Params), C.c_uint32, C.c_void_p]
        L.orc_fk_points.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_rotation.argtypes = [C.c_int, C.c_float, C.c_void_p]
        L.orc_rotation_rpy.argtypes = [C.c_float, C.c_float, C.c_float, C.c_void_p]
        L.orc_fk_model_rexarm.argtypes = [C.c_void_p]
        L.orc_fk_model_points.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_fk_points_as_shipped_matvec.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_grid_index.argtypes = [C.POINTER(Grid), C.c_float, C.c_float, C.c_float]
        L.orc_grid_index.restype = C.c_int64
        L.orc_grid_collides.argtypes = [C.POINTER(Grid), C.c_float, C.c_float, C.c_float, C.c_float]
        L.orc_grid_scatter.argtypes = [C.POINTER(Grid), C.c_void_p, C.c_size_t, C.c_int8, C.c_int]
        L.orc_grid_scatter.restype = C.c_size_t
        L.orc_cost_sample.argtypes = [C.POINTER(Params), C.POINTER(Grid)] + [C.c_void_p] * 9
        L.orc_cost_sample.restype = C.c_float
        L.orc_cost_eval.argtypes = [C.POINTER(Params), C.POINTER(Grid), C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int]
        L.orc_cost_eval.restype = C.c_int64
        L.orc_plan_lerp.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_plan.argtypes = [C.POINTER(Params), C.POINTER(Grid), C.c_void_p, C.c_void_p,
                               C.c_uint32] + [C.c_void_p] * 9 + [C.c_int]
        L.orc_update_vector.argtypes = [C.POINTER(Params)] + [C.c_void_p] * 3
        L.orc_depth_scatter.argtypes = [C.POINTER(Grid), C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                        C.c_void_p, C.c_void_p, C.c_double, C.c_int8, C.c_int]
        L.orc_depth_scatter.restype = C.c_uint64
        L.orc_distance_field.argtypes = [C.POINTER(Grid), C.c_void_p]
        L.orc_distance_field_brute.argtypes = [C.POINTER(Grid), C.c_void_p]
        L.orc_update_sums.argtypes = [C.POINTER(Params), C.c_int32, C.c_void_p, C.c_void_p,
                                      C.c_uint32, C.c_void_p]
        L.orc_update_from_sum.argtypes = [C.c_int64, C.c_int64]
        L.orc_update_from_sum.restype = C.c_float
        L.orc_build_message.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_uint32, C.c_float] + \
            [C.c_void_p] * 3
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


class Oracle:
    """Holds the parameters + grid of one problem; mirrors doapp_b200.capi.Context."""

    def __init__(self, k=1, m=1024, n=64, d=5, n_noise=1, **kw):
        self.p = Params()
        lib().orc_params_init(C.byref(self.p))
        self.p.k, self.p.m, self.p.n, self.p.d, self.p.n_noise = k, m, n, d, n_noise
        self._keep = {}
        for key, val in kw.items():
            if key in ("limit_min", "limit_max", "fk_model"):
                continue
            if key == "global_samples_per_climber":  # the name doapp_config uses
                key = "m_global"
            if not hasattr(self.p, key):
                raise AttributeError(key)
            setattr(self.p, key, val)
        if "limit_min" in kw:
            self.set_limits(kw["limit_min"], kw["limit_max"])
        if kw.get("fk_model") is not None:
            self.set_fk_model(kw["fk_model"])
        self.g = Grid()
        self.g.kind = GRID_NONE
        self.cells = None

    def set_limits(self, lo, hi):
        self._keep["lo"], self._keep["hi"] = _f32(lo), _f32(hi)
        self.p.limit_min = self._keep["lo"].ctypes.data
        self.p.limit_max = self._keep["hi"].ctypes.data

    def set_fk_model(self, model):
        """FkModel -> URDF-correct chain with per-link spheres; None -> the reference's chain."""
        self._keep["fk"] = model
        self.p.fk_model = C.addressof(model) if model is not None else None

    def set_grid(self, cells: np.ndarray, resolution: float, frame=FRAME_CDG):
        """cells indexed [iz, iy, ix] (flat ix + nx*iy + nx*ny*iz), uint8/int8 or float32."""
        assert cells.ndim == 3
        if cells.dtype == np.float32:
            kind = GRID_F32
            self.cells = np.ascontiguousarray(cells)
        else:
            kind = GRID_U8
            self.cells = np.ascontiguousarray(cells.view(np.uint8) if cells.dtype == np.int8
                                              else cells.astype(np.uint8))
        self.g.kind, self.g.frame = kind, frame
        self.g.nz, self.g.ny, self.g.nx = cells.shape
        self.g.resolution = resolution
        self.g.cells = self.cells.ctypes.data

    # -- stage functions ---------------------------------------------------
    def noise(self, iteration=0):
        out = np.empty((self.p.k, self.p.m, self.p.n_noise, self.p.d), np.float32)
        lib().orc_noise_fill(C.byref(self.p), iteration, _p(out))
        return out

    def grid_index(self, xyz):
        xyz = _f32(xyz).reshape(-1, 3)
        return np.array([lib().orc_grid_index(C.byref(self.g), *map(float, q)) for q in xyz],
                        np.int64)

    def grid_collides(self, xyz):
        xyz = _f32(xyz).reshape(-1, 3)
        cd = self.p.collision_distance
        return np.array([lib().orc_grid_collides(C.byref(self.g), float(q[0]), float(q[1]),
                                                 float(q[2]), cd) for q in xyz], np.uint8)

    def scatter(self, xyz, val=1, clear_first=True):
        xyz = _f32(xyz).reshape(-1, 3)
        return lib().orc_grid_scatter(C.byref(self.g), _p(xyz), xyz.shape[0], val, int(clear_first))

    def cost_sample(self, base, start, noise, details=False):
        base, start, noise = _f32(base), _f32(start), _f32(noise)
        n, d = self.p.n, self.p.d
        coll = C.c_int32()
        if not details:
            s = lib().orc_cost_sample(C.byref(self.p), C.byref(self.g), _p(base), _p(start),
                                      _p(noise), C.addressof(coll), None, None, None, None, None)
            return s, coll.value
        noisy, vel, acc = (np.empty((n, d), np.float32) for _ in range(3))
        smooth, viol = C.c_float(), C.c_int32()
        s = lib().orc_cost_sample(C.byref(self.p), C.byref(self.g), _p(base), _p(start), _p(noise),
                                  C.addressof(coll), _p(noisy), _p(vel), _p(acc),
                                  C.addressof(smooth), C.addressof(viol))
        return dict(score=s, collisions=coll.value, noisy=noisy, vel=vel, acc=acc,
                    smooth=smooth.value, violated=viol.value)

    def cost_eval(self, base_traj, start, noise=None, iteration=0, threads=1):
        base_traj, start, noise = _f32(base_traj), _f32(start), _f32(noise)
        S = self.p.k * self.p.m
        scores = np.empty(S, np.float32)
        coll = np.empty(S, np.int32)
        mn = C.c_float()
        am = lib().orc_cost_eval(C.byref(self.p), C.byref(self.g), _p(base_traj), _p(start),
                                 _p(noise), iteration, _p(scores), _p(coll), C.addressof(mn),
                                 threads)
        return dict(scores=scores.reshape(self.p.k, self.p.m),
                    collisions=coll.reshape(self.p.k, self.p.m), argmin=am, min_score=mn.value)

    def plan(self, start, init_traj, iters, noise=None, threads=1):
        start, init_traj, noise = _f32(start), _f32(init_traj), _f32(noise)
        k, n, d = self.p.k, self.p.n, self.p.d
        best_traj, best_vel = np.empty((n, d), np.float32), np.empty((n, d), np.float32)
        best_score, best_climber = C.c_float(), C.c_int64()
        ctraj = np.empty((k, n, d), np.float32)
        cscores, cround = np.empty(k, np.float32), np.empty(k, np.int32)
        fb = np.empty(k, np.uint8)
        lib().orc_plan(C.byref(self.p), C.byref(self.g), _p(start), _p(init_traj), iters, _p(noise),
                       _p(best_traj), _p(best_vel), C.addressof(best_score),
                       C.addressof(best_climber), _p(ctraj), _p(cscores), _p(cround), _p(fb),
                       threads)
        return dict(best_traj=best_traj, best_vel=best_vel, best_score=best_score.value,
                    best_climber=best_climber.value, trajectories=ctraj, scores=cscores,
                    rounded=cround, found_better=fb)

    def depth_scatter(self, depth, intrinsics, extrinsics, crop_size, val=1, clear_first=True) -> int:
        depth = np.ascontiguousarray(depth, np.uint16)
        intr = np.ascontiguousarray(intrinsics, np.float64)
        extr = np.ascontiguousarray(extrinsics, np.float32)
        cams, h, w = depth.shape
        return int(lib().orc_depth_scatter(C.byref(self.g), _p(depth), cams, w, h, _p(intr), _p(extr),
                                           float(crop_size), val, int(clear_first)))

    def distance_field(self, brute=False) -> np.ndarray:
        """Exact Euclidean distance field [nz][ny][nx] f32 of the current U8 grid."""
        out = np.empty((self.g.nz, self.g.ny, self.g.nx), np.float32)
        (lib().orc_distance_field_brute if brute else lib().orc_distance_field)(C.byref(self.g), _p(out))
        return out

    def update_vector(self, scores, noise):
        scores, noise = _f32(scores), _f32(noise)
        out = np.empty((self.p.k, self.p.d), np.float32)
        lib().orc_update_vector(C.byref(self.p), _p(scores), _p(noise), _p(out))
        return out


def update_sums(p_obj, best_rounded, scores, noise):
    """Order-independent integer sums of gt.cu:117-133 for one climber (sums[7] = poisoned)."""
    scores, noise = _f32(scores), _f32(noise)
    sums = np.zeros(8, np.int64)
    lib().orc_update_sums(C.byref(p_obj.p), int(best_rounded), _p(scores), _p(noise),
                          scores.shape[0], _p(sums))
    return sums


def sincos(x):
    x = _f32(np.atleast_1d(x))
    s, c = np.empty_like(x), np.empty_like(x)
    sv, cv = C.c_float(), C.c_float()
    f = lib().orc_sincosf
    for i, v in enumerate(x):
        f(float(v), C.byref(sv), C.byref(cv))
        s[i], c[i] = sv.value, cv.value
    return s, c


def logf(x):
    x = _f32(np.atleast_1d(x))
    return np.array([lib().orc_logf(float(v)) for v in x], np.float32)


def philox(ctr, key):
    ctr = np.ascontiguousarray(ctr, np.uint32)
    key = np.ascontiguousarray(key, np.uint32)
    out = np.empty(4, np.uint32)
    lib().orc_philox4x32_10(_p(ctr), _p(key), _p(out))
    return out


def rotation(axis: int, theta: float) -> np.ndarray:
    """3x3 rotation builder of gt.cu:303-352 (axis 0/1/2 = x/y/z) with the deterministic sin/cos."""
    out = np.empty((3, 3), np.float32)
    lib().orc_rotation(int(axis), float(theta), _p(out))
    return out


def rotation_rpy(roll: float, pitch: float, yaw: float) -> np.ndarray:
    out = np.empty((3, 3), np.float32)
    lib().orc_rotation_rpy(float(roll), float(pitch), float(yaw), _p(out))
    return out


def fk_points(waypoints, as_shipped_matvec=False):
    wp = _f32(waypoints)
    wp = wp.reshape(-1, wp.shape[-1])
    out = np.empty((wp.shape[0], 5, 3), np.float32)
    f = lib().orc_fk_points_as_shipped_matvec if as_shipped_matvec else lib().orc_fk_points
    for i in range(wp.shape[0]):
        f(wp[i].ctypes.data_as(C.c_void_p), wp.shape[1], out[i].ctypes.data_as(C.c_void_p))
    return out


def plan_lerp(start, goal, n):
    start, goal = _f32(start), _f32(goal)
    out = np.empty((n, start.shape[0]), np.float32)
    lib().orc_plan_lerp(_p(start), _p(goal), n, start.shape[0], _p(out))
    return out


def build_message(cur, goal, traj, vel, delta_t=0.1):
    cur, goal, traj, vel = _f32(cur), _f32(goal), _f32(traj), _f32(vel)
    n, d = traj.shape
    pos, velo = np.empty((n + 2, d), np.float32), np.empty((n + 2, d), np.float32)
    tfs = np.empty(n + 2, np.float32)
    lib().orc_build_message(_p(cur), _p(goal), _p(traj), _p(vel), n, d, delta_t, _p(pos), _p(velo),
                            _p(tfs))
    return pos, velo, tfs
