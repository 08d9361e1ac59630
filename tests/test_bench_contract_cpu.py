"""The JSON line bench.py prints, checked on the committed end-of-round lines (profiles/r02_bench.json: own arm, profiles/r02_bench_reference_arm.json:
`--impl reference`): every key of the measurement contract is there with the right type and the derived figures are consistent with each other.  Runs
without a GPU; it guards the contract the driver parses, not the numbers."""
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(name):
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        pytest.skip(f"{name} not committed")
    with open(path) as f:
        return json.loads(f.read().strip().splitlines()[-1])


def test_own_arm_line():
    b = _line("r02_bench.json")
    assert b["metric"] == "nerf_train_samples_per_s" and b["unit"] == "samples/s" and b["higher_is_better"] is True
    assert b["n_gpus"] == 1 and b["scaling"] == "weak" and b["vs_baseline"] is None and b["data"] == "synthetic"
    assert isinstance(b["config"]["workload"], str) and "batch2^18" in b["config"]["workload"] and "model" not in b["config"]
    assert b["steps"] >= 1 and b["warmup"] >= 3 and b["gpu_launches"] > 0
    assert b["value"] > 0 and abs(b["value"] * b["ms_per_step"] * 1e-3 / (1 << 18) - 1.0) < 0.1      # ~2^18 compacted samples per step
    e = b["e2e"]
    assert e["unit"] == b["unit"] and 0 < e["value"] <= b["value"] * 1.05 and e["h2d_bytes_per_step"] == 800 * 800 * 4 and e["d2h_bytes_per_step"] >= 16
    c = b["clocks"]
    assert c["sm_mhz"] > 0 and c["sm_max_mhz"] >= c["sm_mhz"] and isinstance(c["reasons"], list)
    assert not set(c["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    r = b["roofline"]
    assert r["bound"] in ("hbm", "tensor", "issue") and r["unit"] and r["peak"] > 0 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0 < r["frac"] <= 1
    assert "traffic" in r and "kernel" in r
    for name, k in b["kernels"].items():                                                               # every kernel group carries its own bound
        if "frac" in k:
            assert 0 < k["frac"] <= 1.0001 and abs(k["frac"] - k["achieved"] / k["peak"]) < 1e-6, name
    cb = b["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] > 0 and cb["unit"] == b["unit"] and isinstance(cb["sample"], str)


def test_reference_arm_line():
    own, ref = _line("r02_bench.json"), _line("r02_bench_reference_arm.json")
    assert ref["impl"] == "reference" and ref["metric"] == own["metric"] and ref["unit"] == own["unit"] and ref["higher_is_better"] is True
    assert ref["config"]["workload"] == own["config"]["workload"]                                      # the same workload, by name
    assert ref["e2e"] == {"value": ref["value"], "unit": ref["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert ref["cpu_baseline"]["value"] == ref["value"] and ref["cpu_baseline"]["cores"] >= 1 and ref["gpu_launches"] == 0
