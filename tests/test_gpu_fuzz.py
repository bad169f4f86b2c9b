"""Randomised parity: 48 seeded draws from the whole parameter space the C ABI accepts — size (ragged widths and heights,
with and without the TMA path), int16 / uint16, every odd median window 3..15, sharpen mask 3..9 with its sigma and gain,
both sampler border modes, multiply-add order, region-growing mode / connectivity / border expansion, explicit seed lists,
dilation and erosion sizes up to 15 in both structuring-element shapes — each compared with the oracle on every output
(median and sharpened planes bit for bit, masks bit for bit), through the blocking and the ticketed entry point."""
import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu


def _draw(seed):
    rng = np.random.default_rng(1000 + seed)
    w = int(rng.choice([100, 104, 128, 131, 160, 200, 257, 320, 333, 512]))
    h = int(rng.choice([100, 101, 128, 150, 203, 256, 300]))
    kw = dict(
        median_window=int(rng.choice([3, 5, 7, 7, 7, 9, 11, 13, 15])),
        sharpen_mask=int(rng.choice([3, 5, 7, 9, 9])),
        sharpen_sigma=float(rng.choice([0.4, 0.5, 0.8, 1.2])),
        sharpen_gain=float(rng.choice([2.0, 2.0, 0.7, -1.0])),
        sharpen_fma=int(rng.integers(0, 2)),
        border_mode=int(rng.integers(0, 2)),
        rg_mode=int(rng.integers(0, 2)),
        rg_connectivity=int(rng.choice([4, 8])),
        rg_border_expand=int(rng.integers(0, 2)),
        se_shape=int(rng.integers(0, 2)),
        dilate_size=int(rng.choice([3, 3, 5, 7, 11, 15])),
        erode_size=int(rng.choice([0, 3, 5, 9])),
    )
    unsigned = bool(rng.integers(0, 2)) and kw["border_mode"] == 0      # zero border with uint16 is refused by design
    kind = "lowcontrast" if rng.integers(0, 3) == 0 else "default"
    n = int(rng.integers(1, 4))
    raw = np.stack([cm.phantom(w, h, 70 + seed, s, kind=kind, dtype=np.uint16 if unsigned else np.int16) for s in range(n)])
    if rng.integers(0, 4) == 0:   # raw noise over the whole value range instead of a phantom
        lo, hi = (0, 65536) if unsigned else (-32768, 32768)
        raw = rng.integers(lo, hi, raw.shape).astype(raw.dtype)
        kw.update(rg_min=0.9, rg_max=4.0)
    seeds = None
    if rng.integers(0, 3) == 0:
        seeds = np.stack([rng.integers(0, w, 7), rng.integers(0, h, 7)], axis=1).astype(np.int32)
    return raw, kw, seeds


@pytest.mark.parametrize("seed", range(48))
def test_random_parameter_sets_match_the_oracle(lib, seed):
    raw, kw, seeds = _draw(seed)
    want = ["median", "sharpened", "rg_mask", "dilated", "dilated_bits", "rg_iterations"] + (["eroded"] if kw["erode_size"] else [])
    with lib.Context([0]) as ctx:
        got = ctx.process_batch(raw, lib.default_params(**kw), seeds=seeds, want=tuple(want))
        tick = ctx.wait(ctx.process_batch_async(raw, lib.default_params(**kw), seeds=seeds, want=("dilated", "rg_mask")))
        fast = ctx.process_batch(raw, lib.default_params(**kw), seeds=seeds, want=("dilated",))   # fast predicate mode of K1
    w = raw.shape[2]
    for i in range(len(raw)):
        p = cm.oparams(**kw)
        if seeds is None:
            rc, exp = cm.o_pipeline(raw[i], p)
        else:   # the oracle's pipeline takes the adaptive seeds; with an explicit list the stages are chained by hand
            if raw.dtype == np.uint16:
                p.input_unsigned = 1
            f = cm.o_clip(cm.o_normalize(raw[i], p.norm_low, p.norm_high, p.norm_min, p.norm_max), p.clip_min, p.clip_max)
            med = cm.o_median(f, p.median_window, p.border_mode)
            sh = cm.o_sharpen(med, p.sharpen_gain, p.sharpen_sigma, p.sharpen_mask, p.border_mode, p.sharpen_fma)
            rc, rg, _ = cm.o_region_grow(sh, p.rg_min, p.rg_max, seeds, p.rg_mode, p.rg_connectivity, p.rg_border_expand)
            exp = {"median": med, "sharpened": sh, "rg_mask": rg, "dilated": cm.o_dilate(rg, p.dilate_size, p.se_shape)[1]}
            if kw["erode_size"]:
                exp["eroded"] = cm.o_erode(rg, p.erode_size, p.se_shape)[1]
        assert rc == 0
        for name in ("median", "sharpened", "rg_mask", "dilated") + (("eroded",) if kw["erode_size"] else ()):
            np.testing.assert_array_equal(got[name][i], exp[name], err_msg=f"seed {seed} slice {i} {name} {kw}")
        np.testing.assert_array_equal(lib.unpack_bits(got["dilated_bits"][i], w), exp["dilated"], err_msg=f"seed {seed} bits")
        np.testing.assert_array_equal(tick["dilated"][i], exp["dilated"], err_msg=f"seed {seed} async")
        np.testing.assert_array_equal(tick["rg_mask"][i], exp["rg_mask"], err_msg=f"seed {seed} async rg")
        np.testing.assert_array_equal(fast["dilated"][i], exp["dilated"], err_msg=f"seed {seed} fast mode {kw}")
