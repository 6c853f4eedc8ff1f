"""Host-side checks of bench.py's contract with the driver (no GPU): the two arms name the same
workload with the same config keys, the strong-scaled shards cover the job exactly, the CPU legs
stay bounded, and the reference arm prints one well-formed JSON line."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def test_default_workload_is_the_1m_sample_config_strong_scaled():
    w = bench.WORKLOADS["c4s"]
    assert w["m_total"] == 1_048_576 and w["n"] == 128 and w["iters"] >= 20
    for world in (1, 2, 4, 8, 3):
        parts = [bench.shard(w, r, world) for r in range(world)]
        assert sum(p[0] for p in parts) == w["m_total"]
        assert all(p[2] == w["m_total"] for p in parts)
        offs = [p[1] for p in parts]
        assert offs == sorted(offs) and offs[0] == 0
        assert all(parts[i][1] + parts[i][0] == parts[i + 1][1] for i in range(world - 1))
    # weak workloads: fixed work per GPU
    assert bench.shard(bench.WORKLOADS["c2"], 3, 8) == (1024, 3 * 1024, 8 * 1024)


def test_both_arms_name_the_same_config():
    for key in ("c4s", "c2", "c5"):
        for world in (1, 8):
            a = bench.config_of(bench.WORKLOADS[key], world)
            b = bench.config_of(dict(bench.WORKLOADS[key]), world)
            assert a == b and "workload" in a and "model" not in a
    assert bench.config_of(bench.WORKLOADS["c4s"], 1) == bench.config_of(bench.WORKLOADS["c4s"], 8)


def test_cpu_sample_is_bounded():
    for key, w in bench.WORKLOADS.items():
        if key == "c3":
            continue
        m, iters = bench.cpu_sample_shape(w, 8)
        assert 1 <= iters <= w["iters"] and m >= 1
        assert m * w["n"] * iters <= bench.CPU_BUDGET_WPS


def test_algorithmic_bytes_match_the_survey():
    # 5*c + (4d+8)/n -> 5.22 B (n = 128), 5.44 B (n = 64), 5.88 B (n = 32)
    assert abs(bench.alg_bytes_per_wps(128) - 5.21875) < 1e-9
    assert abs(bench.alg_bytes_per_wps(64) - 5.4375) < 1e-9
    assert abs(bench.alg_bytes_per_wps(32) - 5.875) < 1e-9


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--workload", "c2", "--steps", "1", "--warmup", "0", "--no-cpu"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == bench.METRIC and j["unit"] == bench.UNIT
    assert j["config"] == bench.config_of(bench.WORKLOADS["c2"], 1)
    assert j["e2e"]["value"] == j["value"] and j["e2e"]["h2d_bytes_per_step"] == 0
    assert j["cpu_baseline"]["kind"] == "port" and j["cpu_baseline"]["cores"] >= 1
    assert j["higher_is_better"] is True and j["gpu_launches"] == 0
