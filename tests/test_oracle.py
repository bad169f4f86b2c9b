"""Pins the CPU oracle.  The reference ships no golden vectors (parity unpinned), so
the oracle is held to (a) hand-derived known-answer vectors for every integer rule the reference
source states, (b) independent implementations in scipy.ndimage, (c) algebraic properties."""
import numpy as np
import pytest
import scipy.ndimage as ndi

import common as cm


# ---- a7: seeds (main_sequential.cpp:214-241) -------------------------------------------------
def ref_seeds(w, h):
    cx, cy, ox, oy = w // 2, h // 2, w // 8, h // 8
    s = [(cx, cy), (cx + ox, cy), (cx - ox, cy), (cx, cy + oy), (cx, cy - oy)]
    x = w // 4
    while x < w * 3 // 4:
        y = h // 4
        while y < h * 3 // 4:
            s.append((x, y))
            y += h // 10
        x += w // 10
    return np.array(s, np.int32)


@pytest.mark.parametrize("w,h,count", [(256, 256, 41), (512, 512, 41), (1024, 1024, 41), (100, 100, 30), (320, 200, 30)])
def test_seeds_known_answers(w, h, count):
    s = cm.o_make_seeds(w, h)
    assert len(s) == count
    np.testing.assert_array_equal(s, ref_seeds(w, h))


def test_seeds_512_grid_columns():
    s = cm.o_make_seeds(512, 512)
    assert s[:5].tolist() == [[256, 256], [320, 256], [192, 256], [256, 320], [256, 192]]
    assert sorted(set(s[5:, 0].tolist())) == [128, 179, 230, 281, 332, 383]
    # x-major order: first six grid seeds share x = 128
    assert s[5:11, 0].tolist() == [128] * 6


def test_seeds_tiny_image_rejected():
    assert cm.o_make_seeds(9, 200) is None


# ---- a3/a4: normalise + clip -----------------------------------------------------------------
def test_normalize_formula_and_thresholds():
    raw = np.arange(-32768, 32768, dtype=np.int32).astype(np.int16)
    out = cm.o_normalize(raw)
    v = raw.astype(np.float32)
    t = (v - np.float32(0.0)) / np.float32(10000.0)
    exp = t * np.float32(2.0) + np.float32(0.5)
    np.testing.assert_array_equal(out, exp)
    f = cm.o_clip(out)
    assert np.all(np.diff(f) >= 0)  # monotone: the premise of median/normalise commutation (I1)
    # raw <-> normalised anchors: clip floor .68 <-> raw 900, RG window <-> [1200, 2050]
    idx = lambda r: r + 32768
    assert f[idx(899)] == np.float32(0.68) and f[idx(901)] > np.float32(0.68)
    assert abs(f[idx(1200)] - 0.74) < 1e-6 and abs(f[idx(2050)] - 0.91) < 1e-6


def test_normalize_uint16():
    raw = np.array([0, 1, 40000, 65535], np.uint16)
    out = cm.o_normalize(raw)
    exp = (raw.astype(np.float32) / np.float32(10000.0)) * np.float32(2.0) + np.float32(0.5)
    np.testing.assert_array_equal(out, exp)


# ---- a5: median ------------------------------------------------------------------------------
@pytest.mark.parametrize("k", [3, 5, 7])
def test_median_vs_scipy(k):
    rng = np.random.default_rng(k)
    img = rng.integers(0, 50, (37, 53)).astype(np.float32)  # heavy ties
    np.testing.assert_array_equal(cm.o_median(img, k), ndi.median_filter(img, size=k, mode="nearest"))
    np.testing.assert_array_equal(cm.o_median(img, k, border_mode=1),
                                  ndi.median_filter(img, size=k, mode="constant", cval=0.0))


def test_median_commutes_with_normalise_clip():
    """I1: f monotone => median(f(x)) == f(median(x)) bit-exactly."""
    raw = cm.phantom(96, 80, 1, 2)
    f = lambda r: cm.o_clip(cm.o_normalize(r))
    a = cm.o_median(f(raw), 7)
    med_raw = ndi.median_filter(raw, size=7, mode="nearest")
    np.testing.assert_array_equal(a, f(med_raw))


# ---- a6: sharpen -----------------------------------------------------------------------------
def test_gauss_mask_known_answers():
    m = cm.o_gauss_mask(0.5, 9)
    assert abs(m.sum() - 1.0) < 1e-6
    assert abs(m[4, 4] - 0.6186934) < 1e-6
    assert abs(m[4, 5] - 0.0837311) < 1e-6 and abs(m[5, 4] - 0.0837311) < 1e-6
    assert abs(m[5, 5] - 0.0113318) < 1e-6
    assert m[4, 6] == pytest.approx(2.08e-4, rel=2e-2)
    np.testing.assert_array_equal(m, m.T)


def test_sharpen_independent_numpy():
    rng = np.random.default_rng(0)
    img = rng.uniform(0.68, 1.2, (40, 44)).astype(np.float32)
    out = cm.o_sharpen(img)
    mask = cm.o_gauss_mask().astype(np.float64)
    pad = np.pad(img.astype(np.float64), 4, mode="edge")
    blur = np.zeros_like(img, dtype=np.float64)
    for dx in range(-4, 5):
        for dy in range(-4, 5):
            blur += mask[dy + 4, dx + 4] * pad[4 + dy:4 + dy + 40, 4 + dx:4 + dx + 44]
    exp = img + 2.0 * (img - blur)
    np.testing.assert_allclose(out, exp, atol=2e-6)
    # constant image is (almost) a fixed point
    c = np.full((20, 20), 0.8, np.float32)
    np.testing.assert_allclose(cm.o_sharpen(c), c, atol=1e-6)
    # fused variant differs by rounding only
    np.testing.assert_allclose(cm.o_sharpen(img, use_fma=1), out, atol=2e-6)


# ---- a8: region growing ----------------------------------------------------------------------
def _label_fill(pred, seeds, conn):
    st = ndi.generate_binary_structure(2, 1 if conn == 4 else 2)
    lab, _ = ndi.label(pred, structure=st)
    ids = {lab[y, x] for x, y in seeds if lab[y, x] != 0}
    return np.isin(lab, list(ids)).astype(np.uint8) if ids else np.zeros_like(pred, np.uint8)


@pytest.mark.parametrize("conn", [4, 8])
def test_region_grow_vs_scipy_label(conn):
    rng = np.random.default_rng(conn)
    img = ndi.gaussian_filter(rng.uniform(0, 1, (120, 130)), 2.0).astype(np.float32)
    img = (img - img.min()) / (img.max() - img.min())
    lo, hi = 0.45, 0.6
    seeds = cm.o_make_seeds(130, 120)
    rc, out, it = cm.o_region_grow(img, lo, hi, seeds, conn=conn)
    assert rc == 0 and it >= 1
    pred = (img >= np.float32(lo)) & (img <= np.float32(hi))
    np.testing.assert_array_equal(out, _label_fill(pred, seeds, conn))
    assert out.sum() > 0


def test_region_grow_errors_and_modes():
    img = np.full((20, 20), 0.8, np.float32)
    assert cm.o_region_grow(img, 0.7, 0.9, np.zeros((0, 2), np.int32))[0] == -4
    assert cm.o_region_grow(img, 0.7, 0.9, [[20, 3]])[0] == -5
    # isolated in-range seed: opencl mode labels it, host mode does not
    iso = np.zeros((9, 9), np.float32)
    iso[4, 4] = 0.8
    assert cm.o_region_grow(iso, 0.7, 0.9, [[4, 4]], mode=0)[1].sum() == 1
    assert cm.o_region_grow(iso, 0.7, 0.9, [[4, 4]], mode=1)[1].sum() == 0
    # out-of-range seed with in-range neighbours: host mode grows from the neighbours
    img2 = np.full((9, 9), 0.8, np.float32)
    img2[4, 4] = 0.0
    assert cm.o_region_grow(img2, 0.7, 0.9, [[4, 4]], mode=0)[1].sum() == 0
    assert cm.o_region_grow(img2, 0.7, 0.9, [[4, 4]], mode=1)[1].sum() == 80
    # border_expand = 0: growth cannot pass through image-border pixels
    ring = np.zeros((7, 7), np.float32)
    ring[0, :] = 0.8
    full = cm.o_region_grow(ring, 0.7, 0.9, [[3, 0]], border_expand=1)[1]
    stop = cm.o_region_grow(ring, 0.7, 0.9, [[3, 0]], border_expand=0)[1]
    assert full.sum() == 7 and stop.sum() == 1


# ---- a10/a11: morphology ---------------------------------------------------------------------
def _disc(size):
    n = size // 2
    a, b = np.mgrid[-n:n + 1, -n:n + 1]
    return a * a + b * b <= n * n


@pytest.mark.parametrize("size,count", [(3, 5), (5, 13), (7, 29)])
def test_disc_offsets_and_morphology_vs_scipy(size, count):
    assert _disc(size).sum() == count
    one = np.zeros((21, 21), np.uint8)
    one[10, 10] = 1
    rc, d = cm.o_dilate(one, size)
    assert rc == 0 and d.sum() == count
    rng = np.random.default_rng(size)
    m = (ndi.gaussian_filter(rng.uniform(0, 1, (64, 70)), 2.5) > 0.5).astype(np.uint8)
    m[0:5, 0:9] = 1  # touch the border
    np.testing.assert_array_equal(cm.o_dilate(m, size)[1], ndi.binary_dilation(m, _disc(size)).astype(np.uint8))
    np.testing.assert_array_equal(cm.o_erode(m, size)[1],
                                  ndi.binary_erosion(m, _disc(size), border_value=0).astype(np.uint8))
    sq = np.ones((size, size), bool)
    np.testing.assert_array_equal(cm.o_dilate(m, size, 1)[1], ndi.binary_dilation(m, sq).astype(np.uint8))
    # dilate ⊇ mask ⊇ erode
    assert np.all(cm.o_dilate(m, size)[1] >= m) and np.all(m >= cm.o_erode(m, size)[1])


def test_morphology_size_errors():
    m = np.zeros((8, 8), np.uint8)
    assert cm.o_dilate(m, 2)[0] < 0 and cm.o_dilate(m, 1)[0] < 0 and cm.o_erode(m, 4)[0] < 0


# ---- whole pipeline --------------------------------------------------------------------------
def test_pipeline_stage_composition_and_guard():
    raw = cm.phantom(128, 112, 3, 5)
    p = cm.oparams(erode_size=3)
    rc, r = cm.o_pipeline(raw, p)
    assert rc == 0
    f = cm.o_clip(cm.o_normalize(raw))
    med = cm.o_median(f, 7)
    np.testing.assert_array_equal(r["median"], med)
    sh = cm.o_sharpen(med)
    np.testing.assert_array_equal(r["sharpened"], sh)
    _, rg, _ = cm.o_region_grow(sh, 0.74, 0.91, cm.o_make_seeds(128, 112))
    np.testing.assert_array_equal(r["rg_mask"], rg)
    np.testing.assert_array_equal(r["dilated"], cm.o_dilate(rg, 3)[1])
    np.testing.assert_array_equal(r["eroded"], cm.o_erode(rg, 3)[1])
    assert 0 < rg.sum() < rg.size
    assert cm.o_pipeline(cm.phantom(99, 120))[0] == -2  # main_sequential.cpp:189-192


def test_lowcontrast_phantom_covers_half():
    raw = cm.phantom(256, 256, 0, 0, kind="lowcontrast")
    rc, r = cm.o_pipeline(raw)
    assert rc == 0 and r["dilated"].mean() > 0.5


def test_batch_omp_matches_single():
    raws = np.stack([cm.phantom(128, 128, 0, i) for i in range(5)])
    out = np.empty(raws.shape, np.uint8)
    import ctypes as C
    ok = cm.oracle().nm03o_run_batch_omp(raws.ctypes.data_as(C.c_void_p), 5, 128, 128, C.byref(cm.oparams()), 2, 2,
                                         out.ctypes.data_as(C.c_void_p))
    assert ok == 5
    for i in range(5):
        np.testing.assert_array_equal(out[i], cm.o_pipeline(raws[i])[1]["dilated"])


# ---- hand-derived vectors: worked out on paper from SURVEY §8c, not produced by the oracle or by scipy ---------
def test_hand_vector_median_3x3_corners():
    img = np.arange(1, 10, dtype=np.float32).reshape(3, 3)          # 1 2 3 / 4 5 6 / 7 8 9
    # clamp-to-edge, corner (0,0): rows {0,0,1} x cols {0,0,1} -> 1 1 2 | 1 1 2 | 4 4 5 -> sorted 1 1 1 1 [2] 2 4 4 5
    # corner (2,2): 5 6 6 | 8 9 9 | 8 9 9 -> sorted 5 6 6 8 [8] 9 9 9 9;  edge (0,1): 1 2 3 | 1 2 3 | 4 5 6 -> 1 1 2 2 [3] 3 4 5 6
    clamp = cm.o_median(img, 3, 0)
    assert clamp[0, 0] == 2 and clamp[2, 2] == 8 and clamp[0, 1] == 3 and clamp[1, 1] == 5
    # zero border, corner (0,0): 0 0 0 | 0 1 2 | 0 4 5 -> 0 0 0 0 [0] 1 2 4 5;  corner (2,2): 5 6 0 | 8 9 0 | 0 0 0 -> 0 0 0 0 [0] 5 6 8 9
    # edge (1,0): 0 1 2 | 0 4 5 | 0 7 8 -> 0 0 0 1 [2] 4 5 7 8
    zero = cm.o_median(img, 3, 1)
    assert zero[0, 0] == 0 and zero[2, 2] == 0 and zero[1, 0] == 2 and zero[1, 1] == 5


def test_hand_vector_region_growing_5x5():
    v = np.array([[1, 1, 0, 1, 1],
                  [0, 1, 0, 0, 1],
                  [0, 1, 1, 0, 0],
                  [0, 0, 1, 0, 1],
                  [1, 0, 0, 1, 1]], np.float32) * 0.8        # 0.8 is inside [0.74, 0.91], 0 is not
    # 4-connectivity from (x=1, y=1): (1,1) (0,0) (1,0) (1,2) (2,2) (2,3) and nothing else; the diagonal neighbours
    # (3,4)/(4,4)/(4,3) are only reached with 8-connectivity, the top-right block and (0,4) never
    exp4 = np.array([[1, 1, 0, 0, 0],
                     [0, 1, 0, 0, 0],
                     [0, 1, 1, 0, 0],
                     [0, 0, 1, 0, 0],
                     [0, 0, 0, 0, 0]], np.uint8)
    exp8 = exp4.copy()
    exp8[4, 3] = exp8[4, 4] = exp8[3, 4] = 1
    rc, m4, _ = cm.o_region_grow(v, 0.74, 0.91, [(1, 1)], conn=4)
    assert rc == 0
    np.testing.assert_array_equal(m4, exp4)
    rc, m8, _ = cm.o_region_grow(v, 0.74, 0.91, [(1, 1)], conn=8)
    np.testing.assert_array_equal(m8, exp8)
    # a seed on an out-of-range pixel grows nothing in the OpenCL mode; two seeds give the union of their components
    rc, m0, _ = cm.o_region_grow(v, 0.74, 0.91, [(2, 0)], conn=4)
    assert rc == 0 and m0.sum() == 0
    rc, mu, _ = cm.o_region_grow(v, 0.74, 0.91, [(1, 1), (4, 0)], conn=4)
    expu = exp4.copy()
    expu[0, 3] = expu[0, 4] = expu[1, 4] = 1
    np.testing.assert_array_equal(mu, expu)
    # host mode: the seed pixel is only labelled when reached from a neighbour — an isolated in-range seed stays 0
    rc, mh, _ = cm.o_region_grow(v, 0.74, 0.91, [(0, 4)], mode=1, conn=4)
    assert rc == 0 and mh.sum() == 0


def test_hand_vector_single_pixel_morphology():
    m = np.zeros((7, 7), np.uint8)
    m[3, 3] = 1
    plus = np.zeros((7, 7), np.uint8)
    plus[3, 2:5] = 1
    plus[2:5, 3] = 1                                           # size 3 -> N = 1 -> disc a^2+b^2 <= 1: the 5-pixel plus
    rc, d = cm.o_dilate(m, 3, 0)
    np.testing.assert_array_equal(d, plus)
    rc, d = cm.o_dilate(m, 3, 1)                               # square SE: the full 3x3 block
    assert d.sum() == 9 and d[2:5, 2:5].all()
    rc, e = cm.o_erode(plus, 3, 0)                             # the plus eroded by the plus: only its centre survives
    np.testing.assert_array_equal(e, m)
    rc, e = cm.o_erode(plus, 3, 1)
    assert e.sum() == 0                                        # no 3x3 block fits into a plus
    corner = np.zeros((5, 5), np.uint8)
    corner[0:2, 0:2] = 1
    rc, e = cm.o_erode(corner, 3, 0)
    assert e.sum() == 0                                        # outside the image counts as 0: border pixels never survive
    rc, d5 = cm.o_dilate(m, 5, 0)                              # N = 2: 13 pixels (rows of 1, 3, 5, 3, 1)
    assert d5.sum() == 13 and [int(r.sum()) for r in d5[1:6]] == [1, 3, 5, 3, 1]
