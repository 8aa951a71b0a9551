"""CPU suite: the JSON line bench.py printed on the B200 (the last one committed under profiles/) carries every key the measurement
contract names, with consistent values - a schema check of the artefact, not a measurement."""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _last(pattern):
    files = glob.glob(os.path.join(ROOT, "profiles", pattern))
    assert files, pattern
    return max(files, key=lambda f: int(re.search(r"r2_s(\d+)_", os.path.basename(f)).group(1)))


def test_last_bench_line_follows_the_contract():
    d = json.loads(open(_last("r2_s*_bench.json")).readline())
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data",
              "config", "e2e", "gpu_launches", "roofline", "cpu_baseline", "clocks"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["higher_is_better"] is True and d["dtype"] == "f64" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    updates = d["config"]["ntaxa"] - 2
    updates *= d["config"]["patterns"] * d["config"]["ncat"]
    assert abs(d["value"] - updates / (d["ms_per_step"] * 1e-3)) <= 1e-6 * d["value"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == d["config"]["ntaxa"] * d["config"]["patterns"] and e["d2h_bytes_per_step"] > 0
    assert 0 < e["value"] < d["value"], "end to end cannot beat the device-resident number"
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0 < r["frac"] < 1
    assert r["traffic"] and abs(r["traffic"] / r["bytes_per_launch"] - 1) < 0.02, "ncu-measured DRAM traffic vs the bytes the kernel is designed to move"
    c = d["cpu_baseline"]
    assert c["kind"] == "reference" and c["cores"] >= 1 and c["value"] > 0 and "sample" in c
    assert d["gpu_launches"] > 0
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert d["parity"]["ok"] is True and all(x["parity"]["ok"] for x in d["configs"])


def test_last_reference_arm_line_follows_the_contract():
    d = json.loads(open(_last("r2_s*_bench_reference.json")).readline())
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["unit"] == "updates/s"
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["value"] == d["value"]
    assert d["config"]["patterns"] == 1000000 and d["config"]["ntaxa"] == 100
