"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads, and
exports every symbol include/doapp/doapp.h declares; host-only entry points work;
without a CUDA device doapp_create fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "doapp", "doapp.h")).read()
    return sorted(set(re.findall(r"DOAPP_API[^;]*?\b(doapp_\w+)\s*\(", src)))


def test_header_symbols_exported(capi):
    lib = capi.load_library()
    names = _declared()
    assert len(names) >= 30
    for nm in names:
        assert hasattr(lib, nm), f"{nm} declared in doapp.h but not exported"
    assert sorted(capi.ABI_SYMBOLS) == names


def test_header_compiles_as_c():
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        p = os.path.join(td, "t.c")
        open(p, "w").write('#include "doapp/doapp.h"\nint main(void){doapp_config c; (void)c; return 0;}\n')
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I",
                        os.path.join(ROOT, "include"), "-c", p, "-o", os.path.join(td, "t.o")],
                       check=True)


def test_config_defaults_and_struct_size(capi):
    cfg = capi.default_config()
    assert cfg.struct_size == C.sizeof(capi.Config)
    assert (cfg.num_climbers, cfg.samples_per_climber, cfg.num_waypoints, cfg.waypoint_dim) == (1, 1024, 64, 5)
    assert abs(cfg.delta_t - 0.1) < 1e-7 and cfg.smooth_scale == 0.25
    assert abs(cfg.collision_distance - 0.05) < 1e-7 and cfg.seed == 1234
    assert capi.load_library().doapp_version() == 1
    assert capi.load_library().doapp_status_str(1) == b"INVALID_ARG"


def test_plan_lerp_host(capi, oracle_mod):
    s = np.array([0.1, -0.2, 0.3, 0.0, 1.0], np.float32)
    g = np.array([1.0, 1.0, -2.0, 0.5, 1.0], np.float32)
    for n in (1, 7, 64):
        a, b = capi.plan_lerp(s, g, n), oracle_mod.plan_lerp(s, g, n)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        np.testing.assert_allclose(a[-1], g, rtol=1e-6, atol=1e-7)


def test_create_fails_loudly_without_gpu_or_bad_args(capi):
    import torch
    lib = capi.load_library()
    cfg = capi.default_config()
    cfg.waypoint_dim = 7  # invalid regardless of devices
    h = C.c_void_p()
    assert lib.doapp_create(C.byref(cfg), C.byref(h)) == capi.INVALID_ARG
    assert b"waypoint_dim" in lib.doapp_last_error(None)
    if not torch.cuda.is_available():
        try:
            capi.Context(k=1, m=4, n=4)
        except capi.DoappError as e:
            assert e.status == capi.CUDA_ERROR and "no CPU fallback" in e.message
        else:
            raise AssertionError("doapp_create succeeded without a CUDA device")


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "doapp_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, f
                assert "doapp_oracle" not in txt, f
