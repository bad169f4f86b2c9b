"""The host C++17 side on a GPU: our batch driver and — where they were built from
/root/reference in the build container — the reference's OWN three drivers, compiled unchanged
against the FAST-shaped shim, on a synthetic DICOM cohort.  Masks are checked against the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu

HOST = os.path.join(cm.ROOT, "nm03-capstone-project_b200", "host")
DRIVER = os.path.join(HOST, "nm03_driver")
REFBUILD = os.path.join(HOST, "_refbuild")


@pytest.fixture(scope="module")
def cohort(tmp_path_factory):
    if not os.path.exists(DRIVER):
        pytest.fail("nm03_driver is not built (python -c 'import __graft_entry__ as g; g.build()')")
    root = tmp_path_factory.mktemp("data")
    base = root / "Brain-Tumor-Progression" / "T1-Post-Combined-P001-P020"
    subprocess.check_call([DRIVER, "--make-dataset", str(base), "--patients", "3", "--slices", "5", "--size", "256"])
    return root, base


def _host_lib():
    import ctypes as C
    return C.CDLL(os.path.join(HOST, "libnm03_host.so"))


def _phantom(host, w, h, p, s):
    a = np.empty((h, w), np.int16)
    host.nm03h_phantom(w, h, p, s, 0, a.ctypes.data)
    return a


def test_batch_driver_masks_match_oracle(cohort, tmp_path):
    root, base = cohort
    out = tmp_path / "out"
    res = subprocess.run([DRIVER, "--mode", "parallel", "--data", str(base), "--out", str(out), "--dump-masks"],
                         capture_output=True, text=True, check=True)
    summary = json.loads(res.stdout.strip().splitlines()[-1])
    assert summary["ok"] == 15 and summary["patients"] == 3
    assert "Successfully processed 5/5 images" in res.stdout
    host = _host_lib()
    for p in range(3):
        for s in range(5):
            d = out / f"PGBM-{p + 1:03d}"
            mask = np.fromfile(d / f"1-{s + 1:02d}.mask", np.uint8).reshape(256, 256)
            rc, exp = cm.o_pipeline(_phantom(host, 256, 256, p, s))
            np.testing.assert_array_equal(mask, exp["dilated"])
            assert (d / f"1-{s + 1:02d}_original.jpg").stat().st_size > 1000
            assert (d / f"1-{s + 1:02d}_processed.jpg").stat().st_size > 500


def test_sequential_mode_and_test_pipeline(cohort, tmp_path):
    root, base = cohort
    res = subprocess.run([DRIVER, "--mode", "sequential", "--data", str(base), "--out", str(tmp_path / "seq"), "--no-export"],
                         capture_output=True, text=True, check=True)
    assert json.loads(res.stdout.strip().splitlines()[-1])["ok"] == 15
    one = next((base / "PGBM-001").iterdir()) / "1-03.dcm"
    subprocess.run([DRIVER, "--test-pipeline", str(one), "--out", str(tmp_path / "t")], check=True)
    names = sorted(os.listdir(tmp_path / "t"))
    assert names == ["erosion_result.jpg", "final_dilated_result.jpg", "original_image.jpg", "preprocessed_image.jpg",
                     "segmentation.jpg"]


def test_small_slices_are_skipped_not_fatal(tmp_path):
    base = tmp_path / "d"
    subprocess.check_call([DRIVER, "--make-dataset", str(base), "--patients", "1", "--slices", "2", "--size", "96"])
    res = subprocess.run([DRIVER, "--data", str(base), "--out", str(tmp_path / "o"), "--no-export"], capture_output=True, text=True)
    assert res.returncode == 0
    assert "Image dimensions too small: 96x96" in res.stderr          # main_sequential.cpp:189-192
    assert "Successfully processed 0/2 images" in res.stdout


@pytest.mark.parametrize("exe", ["img_processing_sequential", "img_processing_parallel"])
def test_reference_drivers_unchanged_on_the_shim(cohort, tmp_path, exe):
    """The reference's own main_sequential.cpp / main_parallel.cpp, compiled as they are against
    host/compat + libnm03.so in the build container (they cannot be rebuilt on the GPU box)."""
    path = os.path.join(REFBUILD, exe)
    if not os.path.exists(path):
        pytest.skip("reference drivers were not built (no /root/reference at build time)")
    root, base = cohort
    work = tmp_path / "build"
    work.mkdir()
    env = dict(os.environ, NM03_TEST_DATA_PATH=str(root) + "/", OMP_NUM_THREADS="4")
    res = subprocess.run([path], cwd=work, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "Successfully processed 3/3 patients." in res.stdout
    assert res.stdout.count("Successfully processed 5/5 images") == 3
    outdir = tmp_path / ("out-sequential" if "sequential" in exe else "out-parallel")
    jpgs = sorted(os.listdir(outdir / "PGBM-002"))
    assert len(jpgs) == 10 and jpgs[0] == "1-01_original.jpg" and jpgs[1] == "1-01_processed.jpg"
    # the processed JPEG shows the dilated mask: compare its bright area with the oracle's mask area
    from PIL import Image
    host = _host_lib()
    rc, exp = cm.o_pipeline(_phantom(host, 256, 256, 1, 0))
    img = np.asarray(Image.open(outdir / "PGBM-002" / "1-01_processed.jpg").convert("L"))
    area = (img > 100).sum() / 4.0      # 256x256 rendered into a 512x512 canvas
    assert abs(area - exp["dilated"].sum()) / max(exp["dilated"].sum(), 1) < 0.05


def test_driver_shards_patients_over_all_visible_gpus(cohort, tmp_path):
    """Contiguous patient ranges over every GPU of the box (one on the default test box); the masks must not
    depend on the number of GPUs."""
    import torch
    n = torch.cuda.device_count()
    root, base = cohort
    out = tmp_path / "mg"
    res = subprocess.run([DRIVER, "--data", str(base), "--out", str(out), "--dump-masks", "--no-export", "--gpus",
                          ",".join(str(i) for i in range(n))], capture_output=True, text=True, check=True)
    summary = json.loads(res.stdout.strip().splitlines()[-1])
    assert summary["gpus"] == n and summary["ok"] == 15
    host = _host_lib()
    for p in range(3):
        mask = np.fromfile(out / f"PGBM-{p + 1:03d}" / "1-02.mask", np.uint8).reshape(256, 256)
        np.testing.assert_array_equal(mask, cm.o_pipeline(_phantom(host, 256, 256, p, 1))[1]["dilated"])
