"""Host-side C++ helpers without a GPU: DICOM Part-10 writer/reader round trips (explicit and
implicit VR), rejection of unsupported files, the baseline JPEG writer (decoded with PIL), and the
deterministic C++ phantom."""
import ctypes as C
import io
import os
import struct
import subprocess

import numpy as np
import pytest

import common as cm

HOST_DIR = os.path.join(cm.ROOT, "nm03-capstone-project_b200", "host")
HOST_SO = os.path.join(HOST_DIR, "libnm03_host.so")


@pytest.fixture(scope="module")
def host():
    if not os.path.exists(HOST_SO):
        subprocess.check_call(["make", "-C", HOST_DIR, "libnm03_host.so"])
    lib = C.CDLL(HOST_SO)
    lib.nm03h_jpeg_encode.restype = C.c_long
    return lib


def _read(host, path):
    cols, rows, sgn, inst = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    err = C.create_string_buffer(256)
    rc = host.nm03h_dicom_read(path.encode(), C.byref(cols), C.byref(rows), C.byref(sgn), C.byref(inst), None, 0, err, 256)
    if rc:
        raise RuntimeError(err.value.decode())
    px = np.empty((rows.value, cols.value), np.uint16)
    rc = host.nm03h_dicom_read(path.encode(), C.byref(cols), C.byref(rows), C.byref(sgn), C.byref(inst),
                               px.ctypes.data_as(C.c_void_p), C.c_long(px.size), err, 256)
    assert rc == 0
    return px, sgn.value, inst.value


@pytest.mark.parametrize("explicit", [1, 0])
@pytest.mark.parametrize("signed", [1, 0])
def test_dicom_round_trip(host, tmp_path, explicit, signed):
    rng = np.random.default_rng(explicit * 2 + signed)
    img = rng.integers(-2000 if signed else 0, 4096, (131, 203)).astype(np.int16 if signed else np.uint16)
    path = str(tmp_path / "1-07.dcm")
    err = C.create_string_buffer(256)
    assert host.nm03h_dicom_write(path.encode(), img.ctypes.data_as(C.c_void_p), 203, 131, signed, 7, explicit, err, 256) == 0
    raw = open(path, "rb").read()
    assert raw[128:132] == b"DICM"
    px, sgn, inst = _read(host, path)
    assert (sgn, inst) == (signed, 7)
    np.testing.assert_array_equal(px.view(img.dtype), img)


def test_dicom_skips_undefined_length_sequence(host, tmp_path):
    """A private undefined-length sequence with a nested item in front of the pixel data."""
    img = np.arange(100 * 100, dtype=np.uint16).reshape(100, 100)
    path = str(tmp_path / "a.dcm")
    err = C.create_string_buffer(256)
    assert host.nm03h_dicom_write(path.encode(), img.ctypes.data_as(C.c_void_p), 100, 100, 0, 1, 1, err, 256) == 0
    raw = bytearray(open(path, "rb").read())
    pos = raw.find(struct.pack("<HH", 0x7FE0, 0x0010))
    item = struct.pack("<HH2sH", 0x0008, 0x0100, b"SH", 4) + b"ABCD"
    seq = (struct.pack("<HH2sHI", 0x0009, 0x1001, b"SQ", 0, 0xFFFFFFFF) + struct.pack("<HHI", 0xFFFE, 0xE000, 0xFFFFFFFF) + item
           + struct.pack("<HHI", 0xFFFE, 0xE00D, 0) + struct.pack("<HHI", 0xFFFE, 0xE0DD, 0))
    raw[pos:pos] = seq
    open(path, "wb").write(raw)
    px, _, _ = _read(host, path)
    np.testing.assert_array_equal(px, img)


def test_dicom_rejects_bad_files(host, tmp_path):
    p = str(tmp_path / "bad.dcm")
    open(p, "wb").write(b"\0" * 200)
    with pytest.raises(RuntimeError, match="DICM"):
        _read(host, p)
    img = np.zeros((100, 100), np.uint16)
    err = C.create_string_buffer(256)
    host.nm03h_dicom_write(p.encode(), img.ctypes.data_as(C.c_void_p), 100, 100, 0, 1, 1, err, 256)
    raw = open(p, "rb").read().replace(b"1.2.840.10008.1.2.1\0", b"1.2.840.10008.1.2.4.50")  # JPEG baseline
    open(p, "wb").write(raw)
    with pytest.raises(RuntimeError, match="transfer syntax"):
        _read(host, p)
    with pytest.raises(RuntimeError):
        _read(host, str(tmp_path / "missing.dcm"))


def _write_rescaled(host, path, img, signed, slope, intercept):
    err = C.create_string_buffer(256)
    h, w = img.shape
    assert host.nm03h_dicom_write_rescale(path.encode(), img.ctypes.data_as(C.c_void_p), w, h, signed, 1, 1, slope.encode(),
                                          intercept.encode(), err, 256) == 0, err.value


def test_dicom_modality_lut_is_applied(host, tmp_path):
    """RescaleSlope / RescaleIntercept (0028,1053 / 0028,1052): the importer hands on slope * stored + intercept, as DCMTK
    does for FAST's importer; the result is signed when the intercept is negative."""
    rng = np.random.default_rng(5)
    stored = rng.integers(0, 4096, (100, 120)).astype(np.uint16)
    path = str(tmp_path / "ct.dcm")
    _write_rescaled(host, path, stored, 0, "1", "-1024")
    px, sgn, _ = _read(host, path)
    assert sgn == 1
    np.testing.assert_array_equal(px.view(np.int16), stored.astype(np.int32) - 1024)
    _write_rescaled(host, path, stored, 0, "2.0", "10")
    px, sgn, _ = _read(host, path)
    assert sgn == 0
    np.testing.assert_array_equal(px, 2 * stored.astype(np.int32) + 10)
    _write_rescaled(host, path, stored, 0, "1", "0")
    px, sgn, _ = _read(host, path)
    assert sgn == 0
    np.testing.assert_array_equal(px, stored)


def test_dicom_unsupported_rescale_is_refused(host, tmp_path):
    stored = np.full((100, 100), 4000, np.uint16)
    path = str(tmp_path / "r.dcm")
    _write_rescaled(host, path, stored, 0, "0.5", "0")
    with pytest.raises(RuntimeError, match="integer slope and intercept only"):
        _read(host, path)
    _write_rescaled(host, path, stored, 0, "20", "0")          # 80 000 does not fit 16 bits
    with pytest.raises(RuntimeError, match="does not fit 16 bits"):
        _read(host, path)


def _view(host, path):
    cols, rows, off = C.c_int(), C.c_int(), C.c_long()
    err = C.create_string_buffer(256)
    if host.nm03h_dicom_view(path.encode(), C.byref(cols), C.byref(rows), C.byref(off), err, 256):
        raise RuntimeError(err.value.decode())
    return cols.value, rows.value, off.value


@pytest.mark.parametrize("pad", [16384 - 700, 16384 - 640, 16384 - 600, 40000])
def test_positional_reader_rereads_when_the_header_straddles_its_first_block(host, tmp_path, pad):
    """The batch driver parses the first 16 KiB of a file; an element header or an undefined-length sequence cut by that
    boundary must lead to a second, whole-file read — not to a 'truncated' error."""
    img = np.arange(100 * 100, dtype=np.uint16).reshape(100, 100)
    path = str(tmp_path / "a.dcm")
    err = C.create_string_buffer(256)
    assert host.nm03h_dicom_write(path.encode(), img.ctypes.data_as(C.c_void_p), 100, 100, 0, 1, 1, err, 256) == 0
    raw = bytearray(open(path, "rb").read())
    pos = raw.find(struct.pack("<HH", 0x7FE0, 0x0010))
    filler = struct.pack("<HH2sHI", 0x0009, 0x1002, b"OB", 0, pad) + b"\0" * pad       # pushes what follows to the boundary
    item = struct.pack("<HH2sH", 0x0008, 0x0100, b"SH", 4) + b"ABCD"
    seq = (struct.pack("<HH2sHI", 0x0009, 0x1001, b"SQ", 0, 0xFFFFFFFF) + struct.pack("<HHI", 0xFFFE, 0xE000, 0xFFFFFFFF) + item * 6
           + struct.pack("<HHI", 0xFFFE, 0xE00D, 0) + struct.pack("<HHI", 0xFFFE, 0xE0DD, 0))
    raw[pos:pos] = filler + seq
    open(path, "wb").write(raw)
    cols, rows, off = _view(host, path)
    assert (cols, rows) == (100, 100)
    np.testing.assert_array_equal(np.frombuffer(bytes(raw[off:off + 20000]), np.uint16).reshape(100, 100), img)


def test_dicom_nesting_depth_is_bounded(host, tmp_path):
    img = np.zeros((100, 100), np.uint16)
    path = str(tmp_path / "deep.dcm")
    err = C.create_string_buffer(256)
    host.nm03h_dicom_write(path.encode(), img.ctypes.data_as(C.c_void_p), 100, 100, 0, 1, 1, err, 256)
    raw = bytearray(open(path, "rb").read())
    pos = raw.find(struct.pack("<HH", 0x7FE0, 0x0010))
    opener = struct.pack("<HH2sHI", 0x0009, 0x1001, b"SQ", 0, 0xFFFFFFFF) + struct.pack("<HHI", 0xFFFE, 0xE000, 0xFFFFFFFF)
    raw[pos:pos] = opener * 200
    open(path, "wb").write(raw)
    with pytest.raises(RuntimeError, match="nested deeper"):
        _read(host, path)


@pytest.mark.parametrize("channels", [1, 3])
def test_jpeg_decodes_close_to_source(host, channels):
    from PIL import Image
    yy, xx = np.mgrid[0:200, 0:312]
    img = ((np.sin(xx / 17.0) + np.cos(yy / 23.0)) * 60 + 128).astype(np.uint8)
    if channels == 3:
        img = np.stack([img, img[::-1], 255 - img], -1).copy()
    out = np.empty(img.size + 4096, np.uint8)
    n = host.nm03h_jpeg_encode(img.ctypes.data_as(C.c_void_p), 312, 200, channels, 92, out.ctypes.data_as(C.c_void_p),
                               C.c_long(out.size))
    assert n > 0
    dec = np.asarray(Image.open(io.BytesIO(out[:n].tobytes())))
    assert dec.shape == img.shape
    assert np.abs(dec.astype(int) - img.astype(int)).mean() < 2.0


def test_cpp_phantom_is_deterministic_and_in_spec(host):
    a, b = np.empty((512, 512), np.int16), np.empty((512, 512), np.int16)
    host.nm03h_phantom(512, 512, 3, 7, 0, a.ctypes.data_as(C.c_void_p))
    host.nm03h_phantom(512, 512, 3, 7, 0, b.ctypes.data_as(C.c_void_p))
    np.testing.assert_array_equal(a, b)
    assert a.min() >= 0 and a.max() <= 4095
    assert 900 < np.median(a[200:300, 100:160]) < 1100           # tissue
    rc, r = cm.o_pipeline(a)
    assert rc == 0 and 0 < r["dilated"].sum() < 0.3 * a.size     # blobs are segmented, not the whole head
    host.nm03h_phantom(512, 512, 3, 7, 1, b.ctypes.data_as(C.c_void_p))
    assert cm.o_pipeline(b)[1]["dilated"].mean() > 0.5           # low-contrast phantom covers > 50 %


# ---- JPEG writer: byte-for-byte regression vectors ---------------------------------------------------
def _jpeg_cases():
    """Inputs of tests/golden/jpeg_md5.json (hashes of the encoder's output before its bit writer, DCT and
    entropy loop were vectorised; the faster encoder must still produce the same bytes)."""
    rng = np.random.default_rng(7)
    yy, xx = np.mgrid[0:96, 0:128]
    smooth = (127 + 100 * np.sin(xx / 9.0) * np.cos(yy / 7.0)).astype(np.uint8)
    yield "smooth_q90", smooth, 90
    yield "smooth_q35", smooth, 35
    yield "noise_q90", rng.integers(0, 256, (64, 72)), 90
    yield "noise_q100", rng.integers(0, 256, (40, 40)), 100
    yield "extremes", np.where(rng.random((64, 64)) < 0.5, 0, 255), 95     # long codes, 0xFF bytes in the stream
    yield "odd_size", (xx[:37, :51] * 5 + yy[:37, :51] * 3) % 256, 90
    yield "flat", np.full((48, 48), 200), 90
    mask = np.zeros((128, 128), np.uint8)
    mask[30:90, 40:100] = 153
    mask[50:60, 50:60] = 255
    yield "mask", mask, 90
    rgb = np.stack([smooth, np.roll(smooth, 5, 1), 255 - smooth], -1)
    yield "rgb_q90", rgb, 90
    yield "rgb_odd", rgb[:45, :77], 60
    yield "big_noise", rng.normal(128, 40, (512, 512)).clip(0, 255), 90


def _encode(host, img, quality):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape[:2]
    ch = 1 if img.ndim == 2 else img.shape[2]
    out = np.empty(w * h * ch * 2 + 4096, np.uint8)
    n = host.nm03h_jpeg_encode(img.ctypes.data_as(C.c_void_p), w, h, ch, quality, out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
    assert n > 0
    return out[:n].tobytes()


def test_jpeg_bytes_match_pinned_hashes(host):
    import hashlib
    import json
    pinned = json.load(open(os.path.join(cm.ROOT, "tests", "golden", "jpeg_md5.json")))
    seen = set()
    for name, img, q in _jpeg_cases():
        assert hashlib.md5(_encode(host, img, q)).hexdigest() == pinned[name], name
        seen.add(name)
    assert seen == set(pinned)


def test_jpeg_stream_with_stuffed_bytes_decodes(host):
    from PIL import Image
    rng = np.random.default_rng(3)
    img = np.where(rng.random((72, 88)) < 0.5, 0, 255).astype(np.uint8)
    data = _encode(host, img, 95)
    assert b"\xff\x00" in data[600:]                      # the entropy-coded segment needed byte stuffing
    dec = np.asarray(Image.open(io.BytesIO(data)))
    assert dec.shape == img.shape and np.abs(dec.astype(int) - img).mean() < 12


# ---- renderers (CPU restatement of FAST's GL renderers) against numpy/scipy --------------------------
@pytest.mark.parametrize("unsigned", [0, 1])
def test_image_renderer_identity_canvas(host, unsigned):
    rng = np.random.default_rng(11)
    if unsigned:
        img = rng.integers(30000, 42000, (64, 80)).astype(np.uint16)     # crosses 32768
    else:
        img = rng.integers(-900, 3000, (64, 80)).astype(np.int16)
    rgb = np.zeros((64, 80, 3), np.uint8)
    assert host.nm03h_render_image(img.ctypes.data_as(C.c_void_p), 80, 64, unsigned, 80, 64, rgb.ctypes.data_as(C.c_void_p)) == 0
    v = img.astype(np.float32)
    lo, hi = v.min(), v.max()
    g = (v - lo) * (np.float32(255.0) / (hi - lo))
    exp = (np.minimum(np.float32(255), np.maximum(np.float32(0), g)) + np.float32(0.5)).astype(np.uint8)
    for c in range(3):
        np.testing.assert_array_equal(rgb[:, :, c], exp)


def test_image_renderer_fits_a_smaller_slice_into_the_canvas(host):
    img = (np.arange(32 * 48).reshape(32, 48) % 1000).astype(np.int16)
    rgb = np.full((96, 96, 3), 7, np.uint8)
    assert host.nm03h_render_image(img.ctypes.data_as(C.c_void_p), 48, 32, 0, 96, 96, rgb.ctypes.data_as(C.c_void_p)) == 0
    assert (rgb[:15] == 0).all() and (rgb[-15:] == 0).all()      # 48x32 scaled by 2 -> 96x64, centred: black bars
    assert rgb[16:80].max() >= 240 and (rgb[..., 0] == rgb[..., 1]).all()


@pytest.mark.parametrize("radius", [0, 1, 2, 3])
@pytest.mark.parametrize("canvas", [(90, 70), (180, 140), (128, 128)])
def test_segmentation_renderer_border_rule(host, radius, canvas):
    import scipy.ndimage as ndi
    rng = np.random.default_rng(radius * 10 + canvas[0])
    w, h = 90, 70
    lab = (ndi.gaussian_filter(rng.random((h, w)), 2.5) > 0.5).astype(np.uint8)
    lab[0:6, 10:30] = 1                                           # touches the image edge: outside counts as label 0
    lab[40, 50] = 1
    cw, ch = canvas
    rgb = np.zeros((ch, cw, 3), np.uint8)
    assert host.nm03h_render_segmentation(lab.ctypes.data_as(C.c_void_p), w, h, cw, ch, C.c_float(0.6), C.c_float(1.0), radius,
                                          rgb.ctypes.data_as(C.c_void_p)) == 0
    yy, xx = np.mgrid[-radius:radius + 1, -radius:radius + 1]
    disc = (xx * xx + yy * yy) <= radius * radius
    interior = ndi.binary_erosion(lab == 1, structure=disc, border_value=0) if radius > 0 else (lab == 1)
    src = np.zeros((h, w), np.uint8)
    src[lab == 1] = int(np.float32(255.0) * np.float32(1.0) * np.float32(1.0) + np.float32(0.5))         # border: opacity 1
    src[interior] = int(np.float32(0) * np.float32(1 - np.float32(0.6)) + np.float32(255.0) * np.float32(0.6) + np.float32(0.5))
    # aspect-fit, centred, nearest source pixel (the shim's mapping)
    s = min(cw / w, ch / h)
    ox, oy = (cw - w * s) * 0.5, (ch - h * s) * 0.5
    exp = np.zeros((ch, cw), np.uint8)
    for cy in range(ch):
        y = int(np.floor((np.float32(cy) + np.float32(0.5) - np.float32(oy)) / np.float32(s)))
        if not 0 <= y < h:
            continue
        for cx in range(cw):
            x = int(np.floor((np.float32(cx) + np.float32(0.5) - np.float32(ox)) / np.float32(s)))
            if 0 <= x < w:
                exp[cy, cx] = src[y, x]
    for c in range(3):
        np.testing.assert_array_equal(rgb[:, :, c], exp)


# ---- the one-channel canvases the batch driver exports from (fast paths of both renderers) -----------------------
@pytest.mark.parametrize("unsigned", [0, 1])
@pytest.mark.parametrize("shape", [(64, 80), (67, 83), (512, 512)])
def test_image_renderer_grey_canvas_matches_the_formula(host, unsigned, shape):
    """stride-1 identity canvas: eight pixels per step with SSE2; same bytes as the scalar window/level formula."""
    h, w = shape
    rng = np.random.default_rng(w + unsigned)
    img = (rng.integers(30000, 42000, shape).astype(np.uint16) if unsigned else rng.integers(-900, 3000, shape).astype(np.int16))
    grey = np.zeros(shape, np.uint8)
    assert host.nm03h_render_image_grey(img.ctypes.data_as(C.c_void_p), w, h, unsigned, w, h, grey.ctypes.data_as(C.c_void_p)) == 0
    v = img.astype(np.float32)
    lo, hi = v.min(), v.max()
    g = (v - lo) * (np.float32(255.0) / (hi - lo))
    exp = (np.minimum(np.float32(255), np.maximum(np.float32(0), g)) + np.float32(0.5)).astype(np.uint8)
    np.testing.assert_array_equal(grey, exp)
    rgb = np.zeros(shape + (3,), np.uint8)
    assert host.nm03h_render_image(img.ctypes.data_as(C.c_void_p), w, h, unsigned, w, h, rgb.ctypes.data_as(C.c_void_p)) == 0
    np.testing.assert_array_equal(rgb[:, :, 1], grey)


@pytest.mark.parametrize("radius", [0, 1, 2, 3, 5])
@pytest.mark.parametrize("shape", [(70, 90), (64, 64), (100, 131)])
def test_segmentation_renderer_bit_planes(host, radius, shape):
    """Identity canvas: border pixels come from an erosion on bit planes; byte labels and bit-packed labels give the same
    canvas as scipy's erosion with the border disc (outside the image = label 0)."""
    import scipy.ndimage as ndi
    h, w = shape
    rng = np.random.default_rng(radius * 7 + w)
    lab = (ndi.gaussian_filter(rng.random(shape), 2.5) > 0.5).astype(np.uint8)
    lab[0:6, 10:30] = 1
    lab[h - 1, :] = 1
    lab[:, w - 1] = 1
    lab[40, 50] = 1
    yy, xx = np.mgrid[-radius:radius + 1, -radius:radius + 1]
    disc = (xx * xx + yy * yy) <= radius * radius
    interior = ndi.binary_erosion(lab == 1, structure=disc, border_value=0) if radius > 0 else (lab == 1)
    exp = np.zeros(shape, np.uint8)
    exp[lab == 1] = 255
    exp[interior] = int(np.float32(255.0) * np.float32(0.6) + np.float32(0.5))
    wpr = (w + 31) // 32
    padded = np.zeros((h, wpr * 32), np.uint8)
    padded[:, :w] = lab
    bits = np.packbits(padded, axis=1, bitorder="little").view(np.uint32).copy()
    for src, is_bits in ((lab, 0), (bits, 1)):
        grey = np.zeros(shape, np.uint8)
        assert host.nm03h_render_segmentation_grey(src.ctypes.data_as(C.c_void_p), is_bits, w, h, w, h, C.c_float(0.6), C.c_float(1.0),
                                                   radius, grey.ctypes.data_as(C.c_void_p)) == 0
        np.testing.assert_array_equal(grey, exp, err_msg=f"bits={is_bits}")
    # another label value sends the byte image down the general path: label 2 has no colour, label 1 next to it is border
    lab2 = lab.copy()
    lab2[20:24, 20:24] = 2
    grey = np.zeros(shape, np.uint8)
    assert host.nm03h_render_segmentation_grey(lab2.ctypes.data_as(C.c_void_p), 0, w, h, w, h, C.c_float(0.6), C.c_float(1.0), radius,
                                               grey.ctypes.data_as(C.c_void_p)) == 0
    assert (grey[20:24, 20:24] == 0).all()
    interior2 = ndi.binary_erosion(lab2 == 1, structure=disc, border_value=0) if radius > 0 else (lab2 == 1)
    exp2 = np.zeros(shape, np.uint8)
    exp2[lab2 == 1] = 255
    exp2[interior2] = exp[interior].max() if interior.any() else 153
    np.testing.assert_array_equal(grey, exp2)


# ---- blocks transformed elsewhere (nm03_outputs.export_blocks): the entropy coder alone gives the encoder's bytes ---------
@pytest.mark.parametrize("unsigned", [0, 1])
@pytest.mark.parametrize("shape,quality", [((512, 512), 75), ((64, 96), 90), ((128, 72), 30)])
def test_jpeg_from_blocks_equals_direct_encoder(host, unsigned, shape, quality):
    h, w = shape
    host.nm03h_jpeg_from_blocks.restype = C.c_long
    raw = cm.phantom(w, h, 3, 7, dtype=np.uint16 if unsigned else np.int16) if min(w, h) >= 100 else \
        np.random.default_rng(w).integers(0, 3000, shape).astype(np.uint16 if unsigned else np.int16)
    raw[:8, :16] = raw[0, 0]                                       # flat blocks: the direct encoder's DC-only shortcut
    grey = np.zeros(shape, np.uint8)
    assert host.nm03h_render_image_grey(raw.ctypes.data_as(C.c_void_p), w, h, unsigned, w, h, grey.ctypes.data_as(C.c_void_p)) == 0
    direct = np.zeros(w * h * 2 + 4096, np.uint8)
    n1 = host.nm03h_jpeg_encode(grey.ctypes.data_as(C.c_void_p), w, h, 1, quality, direct.ctypes.data_as(C.c_void_p), C.c_long(direct.size))
    nb = (w // 8) * (h // 8)
    blocks, nz = np.zeros((nb, 64), np.int16), np.zeros(nb, np.uint64)
    assert host.nm03h_export_blocks_cpu(raw.ctypes.data_as(C.c_void_p), w, h, unsigned, quality,
                                        blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p)) == 0
    out = np.zeros(direct.size, np.uint8)
    n2 = host.nm03h_jpeg_from_blocks(blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p), w, h, quality,
                                     out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
    assert n1 > 0 and n1 == n2
    np.testing.assert_array_equal(out[:n2], direct[:n1])
    # the map of non-zero terms is what it says
    np.testing.assert_array_equal(nz, (blocks != 0).astype(np.uint64) @ (np.uint64(1) << np.arange(64, dtype=np.uint64)))
    # and the decoded image is the rendering, within JPEG's loss
    from PIL import Image
    dec = np.asarray(Image.open(io.BytesIO(out[:n2].tobytes())))
    assert dec.shape == shape and np.abs(dec.astype(int) - grey.astype(int)).mean() < (8 if min(w, h) >= 100 else 45)   # noise images lose more


def test_jpeg_from_result_blocks_equals_direct_encoder(host):
    import scipy.ndimage as ndi
    host.nm03h_jpeg_from_blocks.restype = C.c_long
    h, w = 256, 320
    rng = np.random.default_rng(9)
    lab = (ndi.gaussian_filter(rng.random((h, w)), 4) > 0.5).astype(np.uint8)
    lab[:5, 30:90] = 1
    grey = np.zeros((h, w), np.uint8)
    assert host.nm03h_render_segmentation_grey(lab.ctypes.data_as(C.c_void_p), 0, w, h, w, h, C.c_float(0.6), C.c_float(1.0), 2,
                                               grey.ctypes.data_as(C.c_void_p)) == 0
    direct = np.zeros(w * h + 4096, np.uint8)
    n1 = host.nm03h_jpeg_encode(grey.ctypes.data_as(C.c_void_p), w, h, 1, 75, direct.ctypes.data_as(C.c_void_p), C.c_long(direct.size))
    nb = (w // 8) * (h // 8)
    blocks, nz, style = np.zeros((nb, 64), np.int16), np.zeros(nb, np.uint64), np.zeros(3, np.uint8)
    assert host.nm03h_export_seg_blocks_cpu(lab.ctypes.data_as(C.c_void_p), 0, w, h, C.c_float(0.6), C.c_float(1.0), 2, 75,
                                            blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p),
                                            style.ctypes.data_as(C.c_void_p)) == 0
    assert style.tolist() == [0, 153, 255] and set(np.unique(grey)) <= {0, 153, 255}
    out = np.zeros(direct.size, np.uint8)
    n2 = host.nm03h_jpeg_from_blocks(blocks.ctypes.data_as(C.c_void_p), nz.ctypes.data_as(C.c_void_p), w, h, 75,
                                     out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
    assert n1 > 0 and n1 == n2
    np.testing.assert_array_equal(out[:n2], direct[:n1])


@pytest.mark.parametrize("radius,opacity,border_opacity", [(2, 0.6, 1.0), (0, 0.5, 0.5), (1, 0.3, 0.9), (3, 0.6, 1.0), (5, 0.8, 0.4)])
@pytest.mark.parametrize("shape", [(512, 512), (104, 136), (64, 64)])
def test_segmentation_jpeg_in_one_step_equals_render_then_encode(host, radius, opacity, border_opacity, shape):
    """encodeSegmentationJpeg (no canvas, flat blocks short-cut) writes the bytes of SegmentationRenderer -> RenderToImage ->
    ImageFileExporter on the same bit-packed mask."""
    import scipy.ndimage as ndi
    host.nm03h_segmentation_jpeg_from_bits.restype = C.c_long
    h, w = shape
    rng = np.random.default_rng(radius * 100 + w)
    lab = (ndi.gaussian_filter(rng.random(shape), 3.0) > 0.5).astype(np.uint8)
    lab[:7, 8:40] = 1                      # a region that runs into the image border
    lab[h - 9:, w - 17:] = 1
    lab[20:44, 16:48] = 1                  # whole blocks of interior pixels
    lab[30, 30] = 0
    wpr = (w + 31) // 32
    padded = np.zeros((h, wpr * 32), np.uint8)
    padded[:, :w] = lab
    bits = np.packbits(padded, axis=1, bitorder="little").view(np.uint32).copy()
    for quality in (75, 40):
        grey = np.zeros(shape, np.uint8)
        assert host.nm03h_render_segmentation_grey(bits.ctypes.data_as(C.c_void_p), 1, w, h, w, h, C.c_float(opacity),
                                                   C.c_float(border_opacity), radius, grey.ctypes.data_as(C.c_void_p)) == 0
        direct = np.zeros(w * h + 4096, np.uint8)
        n1 = host.nm03h_jpeg_encode(grey.ctypes.data_as(C.c_void_p), w, h, 1, quality, direct.ctypes.data_as(C.c_void_p), C.c_long(direct.size))
        out = np.zeros(direct.size, np.uint8)
        n2 = host.nm03h_segmentation_jpeg_from_bits(bits.ctypes.data_as(C.c_void_p), w, h, C.c_float(opacity), C.c_float(border_opacity),
                                                    radius, quality, out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
        assert n1 > 0 and n1 == n2
        np.testing.assert_array_equal(out[:n2], direct[:n1])
    # an empty and a full mask
    for fill in (0, 0xFFFFFFFF):
        b2 = np.full((h, wpr), fill, np.uint32)
        if w % 32:
            b2[:, -1] &= np.uint32((1 << (w % 32)) - 1)
        grey = np.zeros(shape, np.uint8)
        assert host.nm03h_render_segmentation_grey(b2.ctypes.data_as(C.c_void_p), 1, w, h, w, h, C.c_float(opacity), C.c_float(border_opacity),
                                                   radius, grey.ctypes.data_as(C.c_void_p)) == 0
        direct = np.zeros(w * h + 4096, np.uint8)
        n1 = host.nm03h_jpeg_encode(grey.ctypes.data_as(C.c_void_p), w, h, 1, 75, direct.ctypes.data_as(C.c_void_p), C.c_long(direct.size))
        out = np.zeros(direct.size, np.uint8)
        n2 = host.nm03h_segmentation_jpeg_from_bits(b2.ctypes.data_as(C.c_void_p), w, h, C.c_float(opacity), C.c_float(border_opacity),
                                                    radius, 75, out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
        assert n1 == n2 and np.array_equal(out[:n2], direct[:n1])


def test_jpeg_bit_writer_stuffing_on_noise_images(host):
    """Noise at high quality fills the entropy-coded segment with 0xFF bytes (each followed by a stuffed 0x00): the 64-bit
    accumulator of the writer must place them exactly, or PIL's decoder loses the stream."""
    from PIL import Image
    rng = np.random.default_rng(0)
    stuffed = 0
    for q in (100, 95, 75, 30):
        for _ in range(5):
            h, w, ch = int(rng.integers(8, 200)), int(rng.integers(8, 300)), int(rng.choice([1, 3]))
            img = rng.integers(0, 256, (h, w, ch) if ch == 3 else (h, w)).astype(np.uint8)
            out = np.zeros(h * w * ch * 3 + 8192, np.uint8)
            n = host.nm03h_jpeg_encode(img.ctypes.data_as(C.c_void_p), w, h, ch, q, out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
            assert n > 0
            b = out[:n].tobytes()
            stuffed += b[b.index(b"\xff\xda"):].count(b"\xff\x00")
            dec = np.asarray(Image.open(io.BytesIO(b)))
            assert dec.shape == img.shape
            assert np.abs(dec.astype(int) - img.astype(int)).mean() < (4 if q == 100 else 60)
    assert stuffed > 1000


@pytest.mark.parametrize("seed", range(12))
def test_segmentation_jpeg_in_one_step_on_pathological_masks(host, seed):
    """Single pixels, one-pixel lines, checkerboards, full blocks of border pixels, masks that fill whole block rows:
    the short-cuts of encodeSegmentationJpeg must never change a byte."""
    host.nm03h_segmentation_jpeg_from_bits.restype = C.c_long
    rng = np.random.default_rng(1000 + seed)
    h, w = int(rng.choice([64, 96, 128])), int(rng.choice([64, 104, 160]))
    lab = np.zeros((h, w), np.uint8)
    kind = seed % 6
    if kind == 0:
        lab[rng.integers(0, h, 40), rng.integers(0, w, 40)] = 1                      # isolated pixels
    elif kind == 1:
        lab[::3, :] = 1                                                               # one-pixel lines: everything is border
    elif kind == 2:
        lab[(np.add.outer(np.arange(h), np.arange(w)) % 2) == 0] = 1                  # checkerboard
    elif kind == 3:
        lab[8:16, :] = 1                                                              # exactly one block row
        lab[24:48, 8:w - 8] = 1
    elif kind == 4:
        lab[:, :] = 1
        lab[rng.integers(0, h, 5), rng.integers(0, w, 5)] = 0                         # full image with pin holes
    else:
        lab = (rng.random((h, w)) > 0.5).astype(np.uint8)                             # noise
    radius = int(rng.integers(0, 6))
    opacity, border_opacity = float(rng.choice([0.3, 0.6, 1.0])), float(rng.choice([0.5, 1.0]))
    wpr = (w + 31) // 32
    padded = np.zeros((h, wpr * 32), np.uint8)
    padded[:, :w] = lab
    bits = np.packbits(padded, axis=1, bitorder="little").view(np.uint32).copy()
    grey = np.zeros((h, w), np.uint8)
    assert host.nm03h_render_segmentation_grey(bits.ctypes.data_as(C.c_void_p), 1, w, h, w, h, C.c_float(opacity), C.c_float(border_opacity),
                                               radius, grey.ctypes.data_as(C.c_void_p)) == 0
    direct = np.zeros(w * h * 2 + 4096, np.uint8)
    n1 = host.nm03h_jpeg_encode(grey.ctypes.data_as(C.c_void_p), w, h, 1, 75, direct.ctypes.data_as(C.c_void_p), C.c_long(direct.size))
    out = np.zeros(direct.size, np.uint8)
    n2 = host.nm03h_segmentation_jpeg_from_bits(bits.ctypes.data_as(C.c_void_p), w, h, C.c_float(opacity), C.c_float(border_opacity), radius, 75,
                                                out.ctypes.data_as(C.c_void_p), C.c_long(out.size))
    assert n1 > 0 and n1 == n2
    np.testing.assert_array_equal(out[:n2], direct[:n1])
