"""The reference arm of bench.py (`--impl reference`): the CPU oracle in the reference's OpenMP batch
structure, no GPU involved.  Checks the JSON contract on a single short step."""
import json
import os
import subprocess
import sys

import common as cm


def test_reference_arm_prints_one_json_line():
    r = subprocess.run([sys.executable, os.path.join(cm.ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "slices/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert "24 slices 512x512" in d["config"]["workload"]


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(cm.ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_reference_arm_follows_the_config():
    r = subprocess.run([sys.executable, os.path.join(cm.ROOT, "bench.py"), "--impl", "reference", "--config", "c5", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][0])
    assert d["impl"] == "reference" and "low-contrast" in d["config"]["workload"] and d["scaling"] == "weak"
    assert d["value"] > 0 and "24-slice 512x512" in d["cpu_baseline"]["sample"]
