"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  Bit-exact for the median
plane, the predicate, region-growing masks and morphology; the sharpen plane is compared both
bit-exactly (the kernel mirrors the oracle's operation order) and against the stated tolerance
(max abs error <= 1e-5 of the normalised range [0.5, 2.5], i.e. 2e-5)."""
import numpy as np
import pytest
import scipy.ndimage as ndi

import common as cm

pytestmark = pytest.mark.gpu

SHARPEN_TOL = 1e-5 * (2.5 - 0.5)


@pytest.fixture(scope="module")
def ctx(lib):
    c = lib.Context([0])
    yield c
    c.close()


def _cmp_pre(lib, ctx, raw, **kw):
    h, w = raw.shape
    got = ctx.preprocess(raw, lib.default_params(**kw))
    p = cm.oparams(**kw)
    m = p.sharpen_mask
    f = cm.o_clip(cm.o_normalize(raw, p.norm_low, p.norm_high, p.norm_min, p.norm_max), p.clip_min, p.clip_max)
    med = cm.o_median(f, p.median_window, p.border_mode)
    exp = {"median": med,
           "sharpened": cm.o_sharpen(med, p.sharpen_gain, p.sharpen_sigma, m, p.border_mode, p.sharpen_fma)}
    np.testing.assert_array_equal(got["median"], exp["median"])
    assert np.max(np.abs(got["sharpened"] - exp["sharpened"])) <= SHARPEN_TOL
    np.testing.assert_array_equal(got["sharpened"], exp["sharpened"])
    pred = (exp["sharpened"] >= np.float32(p.rg_min)) & (exp["sharpened"] <= np.float32(p.rg_max))
    np.testing.assert_array_equal(lib.unpack_predicate(got["predicate"], w), pred.astype(np.uint8))


@pytest.mark.parametrize("w,h", [(256, 256), (128, 64), (130, 67), (100, 100), (517, 203), (512, 512)])
def test_preprocess_shapes(lib, ctx, w, h):
    _cmp_pre(lib, ctx, cm.phantom(w, h, 1, 3))


@pytest.mark.parametrize("k", [3, 5, 7, 9])
@pytest.mark.parametrize("m", [3, 5, 7, 9])
def test_preprocess_windows(lib, ctx, k, m):
    _cmp_pre(lib, ctx, cm.phantom(160, 96, 2, k + m), median_window=k, sharpen_mask=m, sharpen_sigma=0.8)


def test_preprocess_switches(lib, ctx):
    raw = cm.phantom(200, 150, 4, 1)
    _cmp_pre(lib, ctx, raw, sharpen_fma=1)
    _cmp_pre(lib, ctx, raw, border_mode=1)
    _cmp_pre(lib, ctx, raw, border_mode=1, sharpen_fma=1, median_window=5)


def test_preprocess_full_int16_range_and_ties(lib, ctx):
    rng = np.random.default_rng(0)
    _cmp_pre(lib, ctx, rng.integers(-32768, 32768, (140, 133)).astype(np.int16))
    _cmp_pre(lib, ctx, rng.integers(1195, 1206, (140, 133)).astype(np.int16))   # ties on the RG threshold
    _cmp_pre(lib, ctx, np.full((120, 120), 1200, np.int16))                     # flat image exactly at rg_min
    _cmp_pre(lib, ctx, np.full((120, 120), 2050, np.int16))


def test_preprocess_uint16(lib, ctx):
    rng = np.random.default_rng(1)
    _cmp_pre(lib, ctx, rng.integers(0, 65536, (128, 160)).astype(np.uint16))
    _cmp_pre(lib, ctx, cm.phantom(128, 160, 0, 0, dtype=np.uint16))


# ---- fast predicate mode (no sharpened plane requested): separable blur + exact band recompute ----
def _oracle_pred(raw, **kw):
    p = cm.oparams(**kw)
    if raw.dtype == np.uint16:
        p.input_unsigned = 1
    f = cm.o_clip(cm.o_normalize(raw, p.norm_low, p.norm_high, p.norm_min, p.norm_max), p.clip_min, p.clip_max)
    med = cm.o_median(f, p.median_window, p.border_mode)
    sh = cm.o_sharpen(med, p.sharpen_gain, p.sharpen_sigma, p.sharpen_mask, p.border_mode, p.sharpen_fma)
    return med, ((sh >= np.float32(p.rg_min)) & (sh <= np.float32(p.rg_max))).astype(np.uint8)


def _cmp_fast(lib, ctx, raw, **kw):
    h, w = raw.shape
    med, pred = _oracle_pred(raw, **kw)
    got = ctx.preprocess(raw, lib.default_params(**kw), want=("predicate",))
    np.testing.assert_array_equal(lib.unpack_predicate(got["predicate"], w), pred)
    got = ctx.preprocess(raw, lib.default_params(**kw), want=("median", "predicate"))
    np.testing.assert_array_equal(lib.unpack_predicate(got["predicate"], w), pred)
    np.testing.assert_array_equal(got["median"], med)


@pytest.mark.parametrize("w,h", [(256, 256), (130, 67), (517, 203), (512, 512)])
def test_fast_predicate_shapes(lib, ctx, w, h):
    _cmp_fast(lib, ctx, cm.phantom(w, h, 2, 5))
    _cmp_fast(lib, ctx, cm.phantom(w, h, 2, 6, kind="lowcontrast"))


def test_fast_predicate_adversarial_thresholds(lib, ctx):
    """Images whose sharpened values sit on / next to the two thresholds: the band logic must fall
    back to the exact evaluation exactly where rounding decides the bit."""
    rng = np.random.default_rng(7)
    _cmp_fast(lib, ctx, np.full((120, 140), 1200, np.int16))            # flat at rg_min
    _cmp_fast(lib, ctx, np.full((120, 140), 2050, np.int16))            # flat at rg_max
    _cmp_fast(lib, ctx, rng.integers(1198, 1203, (150, 160)).astype(np.int16))
    _cmp_fast(lib, ctx, rng.integers(2048, 2053, (150, 160)).astype(np.int16))
    steps = np.repeat(np.repeat(rng.integers(1190, 1211, (10, 12)), 16, 0), 16, 1).astype(np.int16)  # flat patches
    _cmp_fast(lib, ctx, steps)
    _cmp_fast(lib, ctx, rng.integers(-32768, 32768, (140, 133)).astype(np.int16))
    _cmp_fast(lib, ctx, rng.integers(0, 65536, (128, 160)).astype(np.uint16))


@pytest.mark.parametrize("kw", [dict(sharpen_fma=1), dict(border_mode=1), dict(median_window=5, sharpen_mask=5),
                                dict(sharpen_sigma=1.0), dict(sharpen_sigma=0.8, sharpen_mask=7, sharpen_gain=-1.5),
                                dict(clip_min=-3.0, norm_low=-1.0, rg_min=-0.2, rg_max=0.1),
                                dict(median_window=3, sharpen_mask=3, sharpen_sigma=0.4),
                                dict(median_window=9), dict(median_window=9, border_mode=1, sharpen_mask=7)])
def test_fast_predicate_switches(lib, ctx, kw):
    _cmp_fast(lib, ctx, cm.phantom(200, 150, 4, 1), **kw)
    rng = np.random.default_rng(3)
    _cmp_fast(lib, ctx, rng.integers(1150, 2100, (130, 131)).astype(np.int16), **kw)


def test_fast_pipeline_masks_only(lib, ctx):
    """Pipeline without the float planes takes the fast K1 mode; masks stay bit-exact."""
    raws = np.stack([cm.phantom(512, 512, 8, i, kind="lowcontrast" if i == 3 else "default") for i in range(4)])
    res = ctx.process_batch(raws, lib.default_params(erode_size=3), want=("dilated", "eroded", "rg_mask"))
    for i in range(4):
        rc, exp = cm.o_pipeline(raws[i], cm.oparams(erode_size=3))
        for name in ("rg_mask", "dilated", "eroded"):
            np.testing.assert_array_equal(res[name][i], exp[name], err_msg=f"slice {i} {name}")


# ---- region growing --------------------------------------------------------------------------
def _rg_image(w, h, seed):
    rng = np.random.default_rng(seed)
    img = ndi.gaussian_filter(rng.uniform(0, 1, (h, w)), 3.0).astype(np.float32)
    return ((img - img.min()) / (img.max() - img.min())).astype(np.float32)


@pytest.mark.parametrize("w,h", [(256, 256), (130, 67), (512, 512), (1024, 1024), (33, 40), (700, 1300)])
@pytest.mark.parametrize("conn", [4, 8])
def test_region_grow_parity(lib, ctx, w, h, conn):
    img = _rg_image(w, h, w + conn)
    seeds = cm.o_make_seeds(w, h)
    for mode in (0, 1):
        for be in (1, 0):
            kw = dict(rg_min=0.42, rg_max=0.58, rg_connectivity=conn, rg_mode=mode, rg_border_expand=be)
            got, it = ctx.region_grow(img, seeds, lib.default_params(**kw))
            rc, exp, _ = cm.o_region_grow(img, 0.42, 0.58, seeds, mode=mode, conn=conn, border_expand=be)
            assert rc == 0 and it >= 1
            np.testing.assert_array_equal(got, exp)


def test_region_grow_serpentine_worst_case(lib, ctx):
    """A one-pixel-wide serpentine: geodesic distance ~ area/2, the hard case for convergence."""
    h = w = 256
    img = np.zeros((h, w), np.float32)
    img[0::2, :] = 0.8
    img[1::4, w - 1] = 0.8
    img[3::4, 0] = 0.8
    seeds = np.array([[0, 0]], np.int32)
    got, it = ctx.region_grow(img, seeds, lib.default_params())
    rc, exp, _ = cm.o_region_grow(img, 0.74, 0.91, seeds)
    np.testing.assert_array_equal(got, exp)
    assert exp.sum() == 128 * 256 + 128 and it > 1


def test_region_grow_errors(lib, ctx):
    img = np.full((64, 64), 0.8, np.float32)
    with pytest.raises(lib.Nm03Error) as e:
        ctx.region_grow(img, np.array([[64, 1]], np.int32))
    assert e.value.code == lib.ERR_SEED_OOB
    with pytest.raises(lib.Nm03Error) as e:
        ctx.region_grow(img, np.zeros((0, 2), np.int32))
    assert e.value.code == lib.ERR_NO_SEEDS


# ---- morphology ------------------------------------------------------------------------------
@pytest.mark.parametrize("w,h", [(256, 256), (130, 67), (512, 512), (1024, 1024), (31, 9)])
@pytest.mark.parametrize("size", [3, 5, 7, 15])
def test_morphology_parity(lib, ctx, w, h, size):
    rng = np.random.default_rng(size)
    m = (ndi.gaussian_filter(rng.uniform(0, 1, (h, w)), 2.0) > 0.5).astype(np.uint8)
    m[: h // 8, : w // 8] = 1
    m[-1, :] = 1
    for se in (0, 1):
        np.testing.assert_array_equal(ctx.morphology(m, size, se, 0), cm.o_dilate(m, size, se)[1])
        np.testing.assert_array_equal(ctx.morphology(m, size, se, 1), cm.o_erode(m, size, se)[1])


def test_morphology_treats_only_one_as_foreground(lib, ctx):
    m = np.zeros((40, 40), np.uint8)
    m[10, 10] = 1
    m[30, 30] = 2   # FAST tests == 1
    np.testing.assert_array_equal(ctx.morphology(m, 3, 0, 0), cm.o_dilate(m, 3)[1])


def test_morphology_errors(lib, ctx):
    m = np.zeros((40, 40), np.uint8)
    for size in (2, 1, 4):
        with pytest.raises(lib.Nm03Error) as e:
            ctx.morphology(m, size)
        assert e.value.code == lib.ERR_PARAM


# ---- whole pipeline --------------------------------------------------------------------------
WANT = ("dilated", "eroded", "rg_mask", "sharpened", "median", "rg_iterations")


def _cmp_pipeline(lib, ctx, raws, **kw):
    res = ctx.process_batch(raws, lib.default_params(**kw), want=WANT)
    for i in range(len(raws)):
        rc, exp = cm.o_pipeline(raws[i], cm.oparams(**kw))
        assert rc == 0
        for name in ("median", "sharpened", "rg_mask", "dilated") + (("eroded",) if kw.get("erode_size", 0) else ()):
            np.testing.assert_array_equal(res[name][i], exp[name], err_msg=f"slice {i} {name}")
        assert res["rg_iterations"][i] >= 1
    return res


def test_pipeline_c1_test_path_256(lib, ctx):
    """BASELINE config 1: one 256x256 slice through all stages incl. the erosion branch."""
    res = _cmp_pipeline(lib, ctx, np.stack([cm.phantom(256, 256, 17, 14)]), erode_size=3)
    assert 0 < res["dilated"].sum() < res["dilated"].size


def test_pipeline_c2_series_512(lib, ctx):
    """BASELINE config 2 (reduced to 6 slices for the oracle's runtime): 512x512 default params."""
    raws = np.stack([cm.phantom(512, 512, 0, i) for i in range(6)])
    res = _cmp_pipeline(lib, ctx, raws)
    assert all(res["dilated"][i].sum() > 0 for i in range(6))


def test_pipeline_c4_large_matrix(lib, ctx):
    """BASELINE config 4 shape: 1024x1024, median window 5, dilation size 7 (radius 3)."""
    _cmp_pipeline(lib, ctx, np.stack([cm.phantom(1024, 1024, 3, i) for i in range(2)]), median_window=5, dilate_size=7)


def test_pipeline_c5_lowcontrast(lib, ctx):
    """BASELINE config 5: grown region covers >50% of each slice."""
    raws = np.stack([cm.phantom(512, 512, 5, i, kind="lowcontrast") for i in range(2)])
    res = _cmp_pipeline(lib, ctx, raws)
    assert res["rg_mask"].mean() > 0.5


def test_pipeline_ragged_and_switches(lib, ctx):
    raws = np.stack([cm.phantom(333, 205, 6, i) for i in range(3)])
    _cmp_pipeline(lib, ctx, raws, erode_size=5, rg_connectivity=8, se_shape=1, sharpen_fma=1)
    _cmp_pipeline(lib, ctx, raws, rg_mode=1, border_mode=1, rg_border_expand=0)


def test_pipeline_median_window_9_and_unsupported_windows(lib, ctx):
    """N4 generality: 9x9 is implemented (slower network); windows above 255 are refused, even ones are
    a parameter error as in FAST."""
    raws = np.stack([cm.phantom(300, 260, 9, i) for i in range(2)])
    _cmp_pipeline(lib, ctx, raws, median_window=9, dilate_size=5)
    for k, code in ((257, lib.ERR_UNSUPPORTED), (4, lib.ERR_PARAM), (1, lib.ERR_PARAM)):
        with pytest.raises(lib.Nm03Error) as e:
            ctx.process_batch(raws, lib.default_params(median_window=k))
        assert e.value.code == code


# ---- median windows 11, 13, 15: the generic rank search (no generated network) ---------------------
@pytest.mark.parametrize("k", [11, 13, 15])
@pytest.mark.parametrize("m", [3, 9])
def test_preprocess_large_windows(lib, ctx, k, m):
    _cmp_pre(lib, ctx, cm.phantom(160, 110, 2, k + m), median_window=k, sharpen_mask=m, sharpen_sigma=0.8)


@pytest.mark.parametrize("k", [11, 15])
def test_preprocess_large_windows_value_ranges_and_borders(lib, ctx, k):
    rng = np.random.default_rng(k)
    _cmp_pre(lib, ctx, rng.integers(-32768, 32768, (140, 133)).astype(np.int16), median_window=k)   # all 16 bits in play
    _cmp_pre(lib, ctx, rng.integers(0, 65536, (128, 160)).astype(np.uint16), median_window=k)
    _cmp_pre(lib, ctx, rng.integers(1195, 1206, (140, 133)).astype(np.int16), median_window=k)      # heavy ties
    _cmp_pre(lib, ctx, np.full((120, 120), 1200, np.int16), median_window=k)                         # one value only
    _cmp_pre(lib, ctx, cm.phantom(200, 150, 4, 1), median_window=k, border_mode=1)                   # zero border sentinel
    _cmp_pre(lib, ctx, cm.phantom(517, 203, 4, 2), median_window=k)                                  # no TMA (width % 8), ragged tiles
    _cmp_fast(lib, ctx, cm.phantom(256, 256, 2, 5), median_window=k)                                 # fast predicate mode


def test_pipeline_median_window_13(lib, ctx):
    raws = np.stack([cm.phantom(512, 512, 9, i) for i in range(3)])
    _cmp_pipeline(lib, ctx, raws, median_window=13, dilate_size=5, erode_size=3)


# ---- median windows above 15: two passes (generic median of the raw samples, then K1 with a 1x1 window) ----------
@pytest.mark.parametrize("k", [17, 21, 33])
def test_preprocess_windows_above_15(lib, ctx, k):
    _cmp_pre(lib, ctx, cm.phantom(160, 110, 2, k), median_window=k, sharpen_mask=3 if k == 21 else 9, sharpen_sigma=0.8)
    _cmp_fast(lib, ctx, cm.phantom(200, 136, 2, k + 1), median_window=k)        # fast predicate mode on the median plane


def test_preprocess_windows_above_15_value_ranges_and_borders(lib, ctx):
    rng = np.random.default_rng(17)
    _cmp_pre(lib, ctx, rng.integers(-32768, 32768, (90, 77)).astype(np.int16), median_window=17)    # all 16 bits in play
    _cmp_pre(lib, ctx, rng.integers(0, 65536, (64, 96)).astype(np.uint16), median_window=19)
    _cmp_pre(lib, ctx, rng.integers(1195, 1206, (70, 66)).astype(np.int16), median_window=23)       # heavy ties
    _cmp_pre(lib, ctx, np.full((40, 50), 1200, np.int16), median_window=17)                          # one value only
    _cmp_pre(lib, ctx, cm.phantom(60, 44, 4, 2), median_window=101)                                  # window larger than the image
    _cmp_pre(lib, ctx, cm.phantom(517, 203, 4, 2), median_window=17)                                 # no TMA (width % 8), ragged tiles
    with pytest.raises(lib.Nm03Error) as e:                                                          # zero border: up to 15 only
        ctx.preprocess(cm.phantom(120, 120, 1, 1), lib.default_params(median_window=17, border_mode=1))
    assert e.value.code == lib.ERR_UNSUPPORTED


def test_pipeline_median_window_25(lib, ctx):
    raws = np.stack([cm.phantom(256, 200, 9, i) for i in range(3)])
    _cmp_pipeline(lib, ctx, raws, median_window=25, dilate_size=5, erode_size=3)
    _cmp_pipeline(lib, ctx, raws.astype(np.uint16), median_window=17, input_unsigned=1)


def test_pipeline_dimension_guard(lib, ctx):
    with pytest.raises(lib.Nm03Error) as e:
        ctx.process_batch(np.zeros((2, 99, 300), np.int16))
    assert e.value.code == lib.ERR_TOO_SMALL
    assert "Image dimensions too small: 300x99" in str(e.value)


def test_pipeline_size_independent_properties_full_c2(lib, ctx):
    """Full 24-slice series: idempotence of the fill, dilate ⊇ mask ⊇ erode, determinism."""
    raws = np.stack([cm.phantom(512, 512, 1, i) for i in range(24)])
    a = ctx.process_batch(raws, lib.default_params(erode_size=3), want=WANT)
    b = ctx.process_batch(raws, lib.default_params(erode_size=3), want=WANT)
    for k in ("dilated", "eroded", "rg_mask"):
        np.testing.assert_array_equal(a[k], b[k])
    assert np.all(a["dilated"] >= a["rg_mask"]) and np.all(a["rg_mask"] >= a["eroded"])
    seeds = lib.make_seeds(512, 512)
    for i in (0, 11, 23):
        pred = ((a["sharpened"][i] >= np.float32(0.74)) & (a["sharpened"][i] <= np.float32(0.91)))
        assert np.all(pred[a["rg_mask"][i] == 1])                       # mask inside the predicate
        again, _ = ctx.region_grow(a["sharpened"][i], seeds)              # fill is idempotent
        np.testing.assert_array_equal(again, a["rg_mask"][i])
        lab, _ = ndi.label(pred)
        ids = {lab[y, x] for x, y in seeds if lab[y, x]}
        np.testing.assert_array_equal(a["rg_mask"][i], np.isin(lab, list(ids)).astype(np.uint8))


# ---- re-entrancy: the reference runs 16 host threads through one device (main_parallel.cpp:336-343,401) ----
def test_concurrent_host_threads_share_one_context(lib, ctx):
    import threading
    raws = [np.stack([cm.phantom(256 + 32 * (t % 3), 224, t, i) for i in range(3)]) for t in range(12)]
    exp = [[cm.o_pipeline(r[i])[1]["dilated"] for i in range(3)] for r in raws]
    got, errs = [None] * len(raws), []

    def work(t):
        try:
            for _ in range(3):   # several rounds so that slots and scratch are reused across threads
                got[t] = ctx.process_batch(raws[t], lib.default_params(), want=("dilated",))["dilated"].copy()
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    ths = [threading.Thread(target=work, args=(t,)) for t in range(len(raws))]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs
    for t in range(len(raws)):
        for i in range(3):
            np.testing.assert_array_equal(got[t][i], exp[t][i], err_msg=f"thread {t} slice {i}")


@pytest.mark.parametrize("w,h", [(100, 290), (512, 512), (300, 1024)])
@pytest.mark.parametrize("conn", [4, 8])
def test_region_grow_from_a_seed_next_to_every_internal_boundary(lib, ctx, w, h, conn):
    """Regression (found by the property test): growth that starts exactly across a row-segment, word-column
    or CTA boundary from the seed.  One-pixel-wide lines, the seed walked along them pixel by pixel."""
    vert = np.zeros((h, w), np.float32)
    vert[20:h - 20, 46] = 0.8
    horiz = np.zeros((h, w), np.float32)
    horiz[h // 2 + 1, 5:w - 5] = 0.8
    diag = np.zeros((h, w), np.float32)
    n = min(w, h) - 10
    diag[np.arange(5, 5 + n), np.arange(5, 5 + n)] = 0.8
    p = lib.default_params(rg_connectivity=conn)
    for img, seeds in ((vert, [(46, y) for y in list(range(20, 60)) + list(range(h // 2 - 20, h // 2 + 20))]),
                       (horiz, [(x, h // 2 + 1) for x in range(25, 70)]),
                       (diag, [(5 + i, 5 + i) for i in range(20, 75)])):
        for sx, sy in seeds:
            got, _ = ctx.region_grow(img, np.array([[sx, sy]], np.int32), p)
            rc, exp, _ = cm.o_region_grow(img, 0.74, 0.91, [[sx, sy]], conn=conn)
            assert np.array_equal(got, exp), f"seed ({sx},{sy}) conn {conn}: got {int(got.sum())} px, expected {int(exp.sum())}"
