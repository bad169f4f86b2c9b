"""Property tests (hypothesis).  CPU: algebraic properties of the oracle's operators.  GPU: the CUDA
pipeline against the oracle on randomly drawn shapes, parameters and data distributions."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import common as cm

import os

GPU_EXAMPLES = int(os.environ.get("NM03_HYP_EXAMPLES", "40"))   # raise for soak runs
SETTINGS = dict(deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.function_scoped_fixture])


def _random_image(rng, w, h, kind):
    if kind == "full":
        return rng.integers(-32768, 32768, (h, w)).astype(np.int16)
    if kind == "ties":
        return rng.integers(1195, 1206, (h, w)).astype(np.int16)
    if kind == "window":
        return rng.integers(800, 2300, (h, w)).astype(np.int16)
    if kind == "blocks":
        b = rng.integers(900, 2200, ((h + 15) // 16, (w + 15) // 16))
        return np.repeat(np.repeat(b, 16, 0), 16, 1)[:h, :w].astype(np.int16)
    return cm.phantom(w, h, int(rng.integers(0, 50)), int(rng.integers(0, 24)))


@settings(max_examples=25, **SETTINGS)
@given(seed=st.integers(0, 2**31), k=st.sampled_from([3, 5, 7, 9]), kind=st.sampled_from(["full", "ties", "window"]))
def test_oracle_median_commutes_with_monotone_map(seed, k, kind):
    """I1: median(f(x)) == f(median(x)) bit-exactly for the monotone f = clip o normalise."""
    import scipy.ndimage as ndi
    rng = np.random.default_rng(seed)
    raw = _random_image(rng, 40 + seed % 17, 33 + seed % 11, kind)
    f = lambda r: cm.o_clip(cm.o_normalize(r))  # noqa: E731
    np.testing.assert_array_equal(cm.o_median(f(raw), k), f(ndi.median_filter(raw, size=k, mode="nearest")))


@settings(max_examples=25, **SETTINGS)
@given(seed=st.integers(0, 2**31), conn=st.sampled_from([4, 8]), size=st.sampled_from([3, 5, 7]))
def test_oracle_fill_and_morphology_properties(seed, conn, size):
    import scipy.ndimage as ndi
    rng = np.random.default_rng(seed)
    img = ndi.gaussian_filter(rng.uniform(0, 1, (60, 70)), 1.5).astype(np.float32)
    lo, hi = np.float32(np.quantile(img, 0.35)), np.float32(np.quantile(img, 0.7))
    seeds = rng.integers(0, [70, 60], (6, 2)).astype(np.int32)
    rc, m, _ = cm.o_region_grow(img, lo, hi, seeds, conn=conn)
    assert rc == 0
    pred = (img >= lo) & (img <= hi)
    assert np.all(pred[m == 1])                                        # mask inside the predicate
    rc, m2, _ = cm.o_region_grow(np.where(m == 1, img, np.float32(-1)), lo, hi, seeds, conn=conn)
    np.testing.assert_array_equal(m2, m)                               # idempotent on its own result
    rc, m_more, _ = cm.o_region_grow(img, lo, hi, np.vstack([seeds, rng.integers(0, [70, 60], (3, 2))]).astype(np.int32), conn=conn)
    assert np.all(m_more >= m)                                         # monotone in the seed set
    d, e = cm.o_dilate(m, size)[1], cm.o_erode(m, size)[1]
    assert np.all(d >= m) and np.all(m >= e)                           # dilate ⊇ mask ⊇ erode
    np.testing.assert_array_equal(cm.o_dilate(m, size)[1], 1 - cm.o_erode(np.pad(1 - m, size, constant_values=1), size)[1][size:-size, size:-size])  # duality


@pytest.mark.gpu
@settings(max_examples=GPU_EXAMPLES, **SETTINGS)
@given(seed=st.integers(0, 2**31), w=st.integers(100, 330), h=st.integers(100, 290),
       k=st.sampled_from([3, 5, 7, 9]), m=st.sampled_from([3, 5, 7, 9]), sigma=st.sampled_from([0.4, 0.5, 0.8, 1.2]),
       conn=st.sampled_from([4, 8]), dil=st.sampled_from([3, 5, 9]), ero=st.sampled_from([0, 3, 7]),
       kind=st.sampled_from(["phantom", "full", "ties", "window", "blocks"]), fma=st.booleans(), want_float=st.booleans(),
       rg_mode=st.sampled_from([0, 0, 1]), border_expand=st.sampled_from([1, 1, 0]), se=st.sampled_from([0, 0, 1]),
       nslices=st.sampled_from([1, 1, 3]))
def test_cuda_pipeline_matches_oracle_on_random_cases(lib, seed, w, h, k, m, sigma, conn, dil, ero, kind, fma, want_float,
                                                      rg_mode, border_expand, se, nslices):
    rng = np.random.default_rng(seed)
    raws = np.stack([_random_image(rng, w, h, kind) for _ in range(nslices)])
    kw = dict(median_window=k, sharpen_mask=m, sharpen_sigma=sigma, rg_connectivity=conn, dilate_size=dil, erode_size=ero,
              sharpen_fma=int(fma), rg_mode=rg_mode, rg_border_expand=border_expand, se_shape=se)
    want = ("dilated", "rg_mask") + (("eroded",) if ero else ()) + (("sharpened", "median") if want_float else ())
    global _CTX
    try:
        _CTX
    except NameError:
        _CTX = lib.Context([0])
    res = _CTX.process_batch(raws, lib.default_params(**kw), want=want)
    for i in range(nslices):
        rc, exp = cm.o_pipeline(raws[i], cm.oparams(**kw))
        assert rc == 0
        for name in want:
            np.testing.assert_array_equal(res[name][i], exp[name], err_msg=f"slice {i} {name}")


@pytest.mark.gpu
@settings(max_examples=max(8, GPU_EXAMPLES // 4), **SETTINGS)
@given(seed=st.integers(0, 2**31), w=st.integers(331, 700), h=st.integers(291, 1100), conn=st.sampled_from([4, 8]),
       kind=st.sampled_from(["phantom", "window", "blocks"]), border_expand=st.sampled_from([1, 0]), nslices=st.sampled_from([1, 2, 5]))
def test_cuda_pipeline_matches_oracle_on_larger_random_slices(lib, seed, w, h, conn, kind, border_expand, nslices):
    """Larger slices exercise other cluster sizes / row splits of the flood fill (default windows, masks only)."""
    rng = np.random.default_rng(seed)
    raws = np.stack([_random_image(rng, w, h, kind) for _ in range(nslices)])
    kw = dict(rg_connectivity=conn, rg_border_expand=border_expand, erode_size=3)
    with lib.Context([0]) as ctx:
        res = ctx.process_batch(raws, lib.default_params(**kw), want=("dilated", "rg_mask", "eroded"))
    for i in range(nslices):
        rc, exp = cm.o_pipeline(raws[i], cm.oparams(**kw))
        assert rc == 0
        for name in ("rg_mask", "dilated", "eroded"):
            np.testing.assert_array_equal(res[name][i], exp[name], err_msg=f"slice {i} {name}")
