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
