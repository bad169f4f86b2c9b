"""Shared test helpers: ctypes binding of the CPU oracle (oracle/libnm03_oracle.so), the product's
C-ABI library loader, and a deterministic numpy phantom.

The oracle is the checker only (see oracle/nm03_oracle.h); the product library never touches it.
"""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ORACLE_SO = os.path.join(ROOT, "oracle", "libnm03_oracle.so")


class OParams(C.Structure):
    _fields_ = [
        ("norm_low", C.c_float), ("norm_high", C.c_float), ("norm_min", C.c_float), ("norm_max", C.c_float),
        ("clip_min", C.c_float), ("clip_max", C.c_float),
        ("median_window", C.c_int),
        ("sharpen_gain", C.c_float), ("sharpen_sigma", C.c_float), ("sharpen_mask", C.c_int),
        ("rg_min", C.c_float), ("rg_max", C.c_float),
        ("dilate_size", C.c_int), ("erode_size", C.c_int),
        ("border_mode", C.c_int), ("sharpen_fma", C.c_int), ("rg_mode", C.c_int),
        ("rg_connectivity", C.c_int), ("rg_border_expand", C.c_int), ("se_shape", C.c_int),
        ("input_unsigned", C.c_int),
    ]


class OOutputs(C.Structure):
    _fields_ = [
        ("median", C.c_void_p), ("sharpened", C.c_void_p), ("rg_mask", C.c_void_p),
        ("dilated", C.c_void_p), ("eroded", C.c_void_p), ("rg_iterations", C.c_void_p),
    ]


_oracle = None


def oracle():
    """Load (building if needed) the oracle shared library."""
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
        _oracle = C.CDLL(ORACLE_SO)
        _oracle.nm03o_make_seeds.restype = C.c_int
        _oracle.nm03o_run_batch_omp.restype = C.c_int
    return _oracle


def oparams(**kw):
    p = OParams()
    oracle().nm03o_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def o_make_seeds(w, h):
    xy = np.zeros((256, 2), np.int32)
    n = oracle().nm03o_make_seeds(w, h, _ptr(xy), 256)
    return xy[:n].copy() if n >= 0 else None


def o_normalize(raw, low=0.5, high=2.5, mn=0.0, mx=10000.0):
    raw = np.ascontiguousarray(raw)
    out = np.empty(raw.shape, np.float32)
    oracle().nm03o_normalize(_ptr(raw), int(raw.dtype == np.uint16), C.c_long(raw.size), C.c_float(low),
                             C.c_float(high), C.c_float(mn), C.c_float(mx), _ptr(out))
    return out


def o_clip(img, lo=0.68, hi=4000.0):
    img = np.ascontiguousarray(img, np.float32)
    out = np.empty_like(img)
    oracle().nm03o_clip(_ptr(img), C.c_long(img.size), C.c_float(lo), C.c_float(hi), _ptr(out))
    return out


def o_median(img, k=7, border_mode=0):
    img = np.ascontiguousarray(img, np.float32)
    h, w = img.shape
    out = np.empty_like(img)
    rc = oracle().nm03o_median(_ptr(img), w, h, k, border_mode, _ptr(out))
    assert rc == 0, rc
    return out


def o_gauss_mask(sigma=0.5, m=9):
    mask = np.empty((m, m), np.float32)
    oracle().nm03o_gauss_mask(C.c_float(sigma), m, _ptr(mask))
    return mask


def o_sharpen(img, gain=2.0, sigma=0.5, m=9, border_mode=0, use_fma=0):
    img = np.ascontiguousarray(img, np.float32)
    h, w = img.shape
    out = np.empty_like(img)
    rc = oracle().nm03o_sharpen(_ptr(img), w, h, C.c_float(gain), C.c_float(sigma), m, border_mode, use_fma, _ptr(out))
    assert rc == 0, rc
    return out


def o_region_grow(img, lo, hi, seeds, mode=0, conn=4, border_expand=1):
    img = np.ascontiguousarray(img, np.float32)
    h, w = img.shape
    seeds = np.ascontiguousarray(seeds, np.int32).reshape(-1, 2)
    out = np.empty((h, w), np.uint8)
    it = C.c_int(0)
    rc = oracle().nm03o_region_grow(_ptr(img), w, h, C.c_float(lo), C.c_float(hi), _ptr(seeds), len(seeds), mode,
                                    conn, border_expand, _ptr(out), C.byref(it))
    return rc, out, it.value


def o_dilate(mask, size=3, se_shape=0):
    mask = np.ascontiguousarray(mask, np.uint8)
    h, w = mask.shape
    out = np.empty_like(mask)
    rc = oracle().nm03o_dilate(_ptr(mask), w, h, size, se_shape, _ptr(out))
    return rc, out


def o_erode(mask, size=3, se_shape=0):
    mask = np.ascontiguousarray(mask, np.uint8)
    h, w = mask.shape
    out = np.empty_like(mask)
    rc = oracle().nm03o_erode(_ptr(mask), w, h, size, se_shape, _ptr(out))
    return rc, out


def o_pipeline(raw, p=None, want=("median", "sharpened", "rg_mask", "dilated", "eroded")):
    """Run the oracle's processSingleImage() restatement on one slice. Returns (rc, dict)."""
    raw = np.ascontiguousarray(raw)
    h, w = raw.shape
    p = p or oparams()
    if raw.dtype == np.uint16:
        p.input_unsigned = 1
    res = {}
    o = OOutputs()
    for name in want:
        if name in ("median", "sharpened"):
            res[name] = np.empty((h, w), np.float32)
        else:
            res[name] = np.empty((h, w), np.uint8)
        setattr(o, name, res[name].ctypes.data)
    it = C.c_int(0)
    o.rg_iterations = C.cast(C.byref(it), C.c_void_p).value
    rc = oracle().nm03o_pipeline(_ptr(raw), w, h, C.byref(p), C.byref(o))
    res["rg_iterations"] = it.value
    if p.erode_size <= 0:
        res.pop("eroded", None)
    return rc, res


def o_batch_omp(slices, p=None, threads=0):
    """The oracle in the reference's OpenMP batch-of-25 structure (main_parallel.cpp:330-347) over a whole
    (n, h, w) series; returns the (n, h, w) uint8 dilated masks.  threads=0 = every core of the box."""
    slices = np.ascontiguousarray(slices)
    n, h, w = slices.shape
    p = p or oparams()
    if slices.dtype == np.uint16:
        p.input_unsigned = 1
    if threads <= 0:
        threads = len(os.sched_getaffinity(0))
    out = np.empty((n, h, w), np.uint8)
    ok = oracle().nm03o_run_batch_omp(_ptr(slices), n, w, h, C.byref(p), 25, threads, _ptr(out))
    assert ok == n, f"oracle processed {ok} of {n} slices"
    return out


def phantom(w, h, patient=0, slice_idx=0, kind="default", dtype=np.int16):
    """Deterministic T1+C-like phantom: 12-bit range, background N(40,15)>=0, head
    ellipse (axes .42w,.48h) tissue N(1000,60), skull ring 600, 1-3 enhancing blobs N(1600,120) near
    the centre (inside the region-growing raw window 1200..2050).  kind='lowcontrast': tissue
    N(1500,150) over the whole ellipse so the grown region covers >50% of the slice."""
    rng = np.random.default_rng((0x4E4D3033 ^ (patient << 16) ^ slice_idx) & 0xFFFFFFFF)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    cx, cy = w / 2.0, h / 2.0
    e = ((xx - cx) / (0.42 * w)) ** 2 + ((yy - cy) / (0.48 * h)) ** 2
    img = np.maximum(rng.normal(40, 15, (h, w)), 0)
    head = e <= 1.0
    if kind == "lowcontrast":
        img[head] = rng.normal(1500, 150, (h, w))[head]
    else:
        img[head] = rng.normal(1000, 60, (h, w))[head]
        ring = (e > 0.88) & head
        img[ring] = rng.normal(600, 30, (h, w))[ring]
        nblob = 1 + (patient + slice_idx) % 3
        for b in range(nblob):
            bx = cx + (b - (nblob - 1) / 2.0) * 0.14 * w + rng.uniform(-0.02, 0.02) * w
            by = cy + rng.uniform(-0.06, 0.06) * h
            r = (0.06 + 0.02 * ((slice_idx % 8) / 8.0)) * min(w, h) * (1.0 + 0.3 * b)
            blob = (xx - bx) ** 2 + (yy - by) ** 2 <= r * r
            img[blob] = rng.normal(1600, 120, (h, w))[blob]
    img = np.clip(np.rint(img), 0, 4095)
    return img.astype(dtype)


def load_product():
    """Import the product package (hyphenated directory name, hence importlib)."""
    return importlib.import_module("nm03-capstone-project_b200")
