"""doapp_b200 — B200-native perturbator hot path (cost evaluation, planner loop,
voxel scatter) behind the C ABI of include/doapp/doapp.h.

The package holds only what the path needs: csrc/ (CUDA kernels + C ABI),
host/ (C++ mirror of the reference's PlanningState) and capi.py (ctypes
binding used by the tests and the bench).  There is no CPU fallback.
"""
from .capi import Context, DoappError, load_library, plan_lerp  # noqa: F401
