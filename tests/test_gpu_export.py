"""nm03_outputs.export_blocks on a GPU: the quantised JPEG blocks of the original's grey rendering must be those of the
host writer (render + transform + quantise + zigzag on the CPU), coefficient for coefficient, and the files the batch
driver writes from them must be byte-identical to the files of its all-CPU export."""
import ctypes as C
import hashlib
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


@pytest.fixture(scope="module")
def host():
    h = C.CDLL(os.path.join(HOST, "libnm03_host.so"))
    h.nm03h_jpeg_from_blocks.restype = C.c_long
    h.nm03h_jpeg_encode.restype = C.c_long
    return h


def _rq(host, quality):
    rq = np.zeros(64, np.float32)
    host.nm03h_jpeg_rq(quality, rq.ctypes.data_as(C.c_void_p))
    return rq


def _cpu_blocks(host, raw, quality):
    h, w = raw.shape
    nb = (h // 8) * (w // 8)
    blocks, nz = np.zeros((nb, 64), np.int16), np.zeros(nb, np.uint64)
    assert host.nm03h_export_blocks_cpu(raw.ctypes.data_as(C.c_void_p), w, h, int(raw.dtype == np.uint16), quality,
                                        blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p)) == 0
    return blocks, nz


@pytest.mark.parametrize("quality", [75, 92, 20])
def test_blocks_equal_the_host_writer(lib, ctx, host, quality):
    raws = np.stack([cm.phantom(512, 512, 3, i) for i in range(5)])
    raws[4, :64, :128] = 777                                            # flat blocks
    res = ctx.process_batch(raws, want=("dilated", "export"), out={"export_rq": _rq(host, quality)})
    for i in range(len(raws)):
        blocks, nz = _cpu_blocks(host, raws[i], quality)
        np.testing.assert_array_equal(res["export_blocks"][i], blocks, err_msg=f"slice {i}")
        np.testing.assert_array_equal(res["export_nz"][i], nz, err_msg=f"slice {i}")
    np.testing.assert_array_equal(res["dilated"][0], cm.o_pipeline(raws[0])[1]["dilated"])   # the pipeline's result is unaffected


def test_blocks_value_ranges_sizes_and_async(lib, ctx, host):
    rng = np.random.default_rng(5)
    cases = [rng.integers(-32768, 32768, (3, 104, 136)).astype(np.int16),           # all 16 bits, ragged block grid
             rng.integers(0, 65536, (2, 128, 160)).astype(np.uint16),
             np.full((2, 112, 104), 1200, np.int16),                                 # one value: empty window, inv = 0
             np.stack([cm.phantom(256, 200, 1, i) for i in range(3)])]
    for raws in cases:
        out = {"export_rq": _rq(host, 75), "dilated_bits": np.empty((len(raws), raws.shape[1], (raws.shape[2] + 31) // 32), np.uint32)}
        n, h, w = raws.shape
        out["export_blocks"] = np.empty((n, (h // 8) * (w // 8), 64), np.int16)
        out["export_nz"] = np.empty((n, (h // 8) * (w // 8)), np.uint64)
        t = ctx.process_batch_async(raws, out=out, want=("dilated_bits", "export"))
        res = ctx.wait(t)
        for i in range(n):
            blocks, nz = _cpu_blocks(host, raws[i], 75)
            np.testing.assert_array_equal(res["export_blocks"][i], blocks)
            np.testing.assert_array_equal(res["export_nz"][i], nz)


@pytest.mark.parametrize("radius,opacity,border_opacity", [(2, 0.6, 1.0), (0, 0.5, 0.5), (1, 0.3, 0.9), (3, 0.6, 1.0), (7, 0.8, 0.4)])
def test_result_blocks_equal_the_host_renderer(lib, ctx, host, radius, opacity, border_opacity):
    """export_seg_blocks: SegmentationRenderer's rendering of the dilated mask, transformed and quantised on the GPU."""
    raws = np.stack([cm.phantom(512, 512, 2, i) for i in range(3)] + [cm.phantom(512, 512, 5, 1, kind="lowcontrast")])
    n, h, w = raws.shape
    nb = (h // 8) * (w // 8)
    style3 = np.zeros(3, np.uint8)
    exp = []
    quality = 75
    # the greys of the style come from the host renderer's own arithmetic
    dummy_b, dummy_n = np.zeros((nb, 64), np.int16), np.zeros(nb, np.uint64)
    assert host.nm03h_export_seg_blocks_cpu(np.zeros((h, w), np.uint8).ctypes.data_as(C.c_void_p), 0, w, h, C.c_float(opacity),
                                            C.c_float(border_opacity), radius, quality, dummy_b.ctypes.data_as(C.c_void_p),
                                            dummy_n.ctypes.data_as(C.c_void_p), style3.ctypes.data_as(C.c_void_p)) == 0
    style = lib.ExportStyle(int(style3[0]), int(style3[1]), int(style3[2]), radius)
    res = ctx.process_batch(raws, lib.default_params(dilate_size=5), want=("dilated_bits", "export"),
                            out={"export_rq": _rq(host, quality), "export_seg_style": style})
    assert res["dilated_bits"].any()
    for i in range(n):
        blocks, nz = np.zeros((nb, 64), np.int16), np.zeros(nb, np.uint64)
        bits = np.ascontiguousarray(res["dilated_bits"][i])
        assert host.nm03h_export_seg_blocks_cpu(bits.ctypes.data_as(C.c_void_p), 1, w, h, C.c_float(opacity), C.c_float(border_opacity),
                                                radius, quality, blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p),
                                                None) == 0
        np.testing.assert_array_equal(res["export_seg_blocks"][i], blocks, err_msg=f"slice {i}")
        np.testing.assert_array_equal(res["export_seg_nz"][i], nz, err_msg=f"slice {i}")
        ob, on = _cpu_blocks(host, raws[i], quality)                      # the original's blocks come with them
        np.testing.assert_array_equal(res["export_blocks"][i], ob)


def test_result_blocks_masks_at_the_image_border(lib, ctx, host):
    """A grown region that runs into the image border (everything in range): border pixels along the edge, ragged sizes."""
    raws = np.full((2, 104, 136), 1500, np.int16)                        # every pixel inside the region-growing window
    raws[1, 40:60, 50:90] = 0
    n, h, w = raws.shape
    nb = (h // 8) * (w // 8)
    style = lib.ExportStyle(0, 153, 255, 2)
    res = ctx.process_batch(raws, want=("dilated_bits", "export"), out={"export_rq": _rq(host, 75), "export_seg_style": style})
    assert lib.unpack_bits(res["dilated_bits"], w).mean() > 0.5
    for i in range(n):
        blocks, nz = np.zeros((nb, 64), np.int16), np.zeros(nb, np.uint64)
        bits = np.ascontiguousarray(res["dilated_bits"][i])
        assert host.nm03h_export_seg_blocks_cpu(bits.ctypes.data_as(C.c_void_p), 1, w, h, C.c_float(0.6), C.c_float(1.0), 2, 75,
                                                blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p), None) == 0
        np.testing.assert_array_equal(res["export_seg_blocks"][i], blocks)
        np.testing.assert_array_equal(res["export_seg_nz"][i], nz)


def test_blocks_need_whole_blocks_and_all_three_pointers(lib, ctx, host):
    with pytest.raises(lib.Nm03Error) as e:
        ctx.process_batch(np.zeros((1, 100, 132), np.int16), want=("dilated", "export"), out={"export_rq": _rq(host, 75)})
    assert e.value.code == lib.ERR_UNSUPPORTED


def test_driver_files_identical_with_and_without_gpu_blocks(tmp_path):
    base = tmp_path / "data" / "Brain-Tumor-Progression" / "T1-Post-Combined-P001-P020"
    subprocess.check_call([DRIVER, "--make-dataset", str(base), "--patients", "2", "--slices", "6", "--size", "512"])
    outs = {}
    for name, extra in (("gpu", []), ("gpu_seg", ["--gpu-seg"]), ("cpu", ["--cpu-dct"])):
        out = tmp_path / name
        subprocess.run([DRIVER, "--mode", "parallel", "--data", str(base), "--out", str(out)] + extra, capture_output=True, text=True, check=True)
        outs[name] = {str(p.relative_to(out)): hashlib.md5(p.read_bytes()).hexdigest() for p in sorted(out.rglob("*.jpg"))}
    assert len(outs["gpu"]) == 24 and outs["gpu"] == outs["cpu"] and outs["gpu_seg"] == outs["cpu"]
