"""Bit-exact parity at BASELINE.json's full configurations: every dilated mask of C2 (one 24-slice 512x512
series), C3 (the 20-patient x 24-slice cohort through the C++ batch driver, patients sharded over every visible
GPU), C4 (1024x1024, median window 5, dilation size 7) and C5 (low-contrast phantom, grown region > 50 % of the
slice) is compared with the oracle run in the reference's OpenMP batch structure (main_parallel.cpp:330-347).
C4 and C5 use a sample of the configuration's slices (the configurations themselves are thousands of
independent slices of the same kind); C2 and C3 are complete."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu

HOST = os.path.join(cm.ROOT, "nm03-capstone-project_b200", "host")
DRIVER = os.path.join(HOST, "nm03_driver")


@pytest.fixture(scope="module")
def ctx(lib):
    c = lib.Context([0])
    yield c
    c.close()


def _host_phantom(w, h, p, s, kind=0):
    host = C.CDLL(os.path.join(HOST, "libnm03_host.so"))
    a = np.empty((h, w), np.int16)
    host.nm03h_phantom(w, h, p, s, kind, a.ctypes.data)
    return a


def test_c2_full_series_512(lib, ctx):
    raw = np.stack([cm.phantom(512, 512, 3, s) for s in range(24)])
    got = ctx.process_batch(raw, lib.default_params(), want=("dilated", "rg_iterations"))
    exp = cm.o_batch_omp(raw)
    np.testing.assert_array_equal(got["dilated"], exp)
    assert exp.sum() > 0 and (got["status"] == 0).all()


def test_c3_cohort_through_the_driver_on_all_gpus(tmp_path):
    import torch
    n_gpus = torch.cuda.device_count()
    base = tmp_path / "d"
    subprocess.check_call([DRIVER, "--make-dataset", str(base), "--patients", "20", "--slices", "24", "--size", "512"],
                          stdout=subprocess.DEVNULL)
    out = tmp_path / "o"
    res = subprocess.run([DRIVER, "--data", str(base), "--out", str(out), "--dump-masks", "--no-export", "--gpus",
                          ",".join(str(i) for i in range(n_gpus))], capture_output=True, text=True, check=True)
    summary = json.loads(res.stdout.strip().splitlines()[-1])
    assert summary["ok"] == 480 and summary["patients"] == 20 and summary["gpus"] == n_gpus
    mismatched = 0
    for p in range(20):
        raw = np.stack([_host_phantom(512, 512, p, s) for s in range(24)])
        exp = cm.o_batch_omp(raw)
        for s in range(24):
            mask = np.fromfile(out / f"PGBM-{p + 1:03d}" / f"1-{s + 1:02d}.mask", np.uint8).reshape(512, 512)
            mismatched += int(not np.array_equal(mask, exp[s]))
    assert mismatched == 0, f"{mismatched} of 480 masks differ from the oracle"


def test_c4_large_matrix_1024_median5_dilation7(lib, ctx):
    raw = np.stack([cm.phantom(1024, 1024, 5, s) for s in range(8)])
    kw = dict(median_window=5, dilate_size=7)
    got = ctx.process_batch(raw, lib.default_params(**kw), want=("dilated",))
    exp = cm.o_batch_omp(raw, cm.oparams(**kw))
    np.testing.assert_array_equal(got["dilated"], exp)
    assert exp.sum() > 0


def test_c5_low_contrast_worst_case_region_growing(lib, ctx):
    raw = np.stack([cm.phantom(512, 512, 7, s, kind="lowcontrast") for s in range(24)])
    got = ctx.process_batch(raw, lib.default_params(), want=("dilated", "rg_iterations"))
    exp = cm.o_batch_omp(raw)
    np.testing.assert_array_equal(got["dilated"], exp)
    assert exp.mean() > 0.5, "the grown region must cover more than half of every slice"
    assert got["rg_iterations"].min() >= 1
