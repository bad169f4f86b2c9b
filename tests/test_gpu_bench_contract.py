"""bench.py on a GPU: one JSON line with every key the measurement contract names, on a short run."""
import json
import os
import subprocess
import sys

import pytest

import common as cm

pytestmark = pytest.mark.gpu


def test_bench_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(cm.ROOT, "bench.py"), "--steps", "24", "--warmup", "3", "--cpu-seconds", "0.3", "--no-io"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "cpu_baseline", "parity_checked", "regions_ms"):
        assert key in d, key
    assert d["n_gpus"] == 1 and d["steps"] == 24 and d["unit"] == "slices/s" and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["value"] > 1000 and d["ms_per_step"] > 0
    assert d["gpu_launches"] == 2 * d["steps"]                       # K1 + K23 per step
    e = d["e2e"]
    assert e["value"] > 1000 and e["h2d_bytes_per_step"] == 24 * 512 * 512 * 2 and e["d2h_bytes_per_step"] == 24 * 512 * 512
    assert e["value"] < d["value"]                                   # host copies inside the timed region
    rf = d["roofline"]
    assert rf["bound"] in ("hbm", "tensor") and rf["unit"] == "GB/s" and rf["peak"] > 1000
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9 and rf["avg_launch_ms"] > 0
    assert rf["bytes_per_launch"] == 24 * 512 * 512 * (2 + 1 / 8)
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] > 0 and cb["unit"] == "slices/s"
    assert "workload" in d["config"] and "model" not in d["config"]
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    assert d["parity_checked"] == dict(d["parity_checked"], slices=48, mismatches=0)
    assert len(d["regions_ms"]) == 5 and abs(min(d["regions_ms"]) / d["steps"] - d["ms_per_step"]) < 1e-9
    assert e["link_ceiling_gbs"] > 10 and 0 < e["frac_of_link"] < 1.5
    alu = rf["alu"]
    assert 0 < alu["frac"] < 1 and alu["ops_per_run"] > 1000 and alu["minmax_lane_ops_per_launch"] > 1e8
    assert rf["frac_headline"] > 0 and rf["kernel_sources_sha"]


@pytest.mark.parametrize("config,scaling", [("c3", "strong"), ("c4", "weak"), ("c5", "weak")])
def test_bench_other_baseline_configs(config, scaling):
    r = subprocess.run([sys.executable, os.path.join(cm.ROOT, "bench.py"), "--config", config, "--steps", "4", "--warmup", "3",
                        "--cpu-seconds", "0.3", "--no-io"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][0])
    assert d["config"]["name"] == config and d["scaling"] == scaling and d["parity_checked"]["mismatches"] == 0
    assert d["value"] > 1000 and d["e2e"]["value"] > 1000
    if config == "c5":
        assert d["roofline"]["region_growing_passes"]["grown_fraction_of_slice"] > 0.5
