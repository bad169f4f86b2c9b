"""The library's alternative code paths, each forced by its environment switch in a fresh process (the switches are
read when the library is loaded): plain loads instead of TMA, no clip-floor skip list, no central-tiles-first
order, static instead of counter-based work distribution, the exact sharpen phase instead of the fast one, a
fixed cluster size, and all of them together.  Every variant must give the oracle's masks bit for bit."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu

CODE = textwrap.dedent("""
    import sys, numpy as np
    sys.path.insert(0, %r); sys.path.insert(0, %r)
    import common as cm
    nm03 = cm.load_product()
    with nm03.Context([0]) as ctx:
        cases = [(np.stack([cm.phantom(512, 512, 6, i, kind="lowcontrast" if i == 1 else "default") for i in range(3)]), dict(erode_size=3)),
                 (np.stack([cm.phantom(136, 104, 2, i) for i in range(2)]), dict(median_window=5, dilate_size=5)),
                 (np.stack([cm.phantom(131, 203, 3, i) for i in range(2)]), dict(rg_connectivity=8))]
        for raws, kw in cases:
            res = ctx.process_batch(raws, nm03.default_params(**kw), want=("dilated", "rg_mask", "eroded", "median"))
            for i in range(raws.shape[0]):
                rc, exp = cm.o_pipeline(raws[i], cm.oparams(**kw))
                assert rc == 0
                np.testing.assert_array_equal(res["median"][i], exp["median"])
                np.testing.assert_array_equal(res["rg_mask"][i], exp["rg_mask"])
                np.testing.assert_array_equal(res["dilated"][i], exp["dilated"])
                if kw.get("erode_size"):
                    np.testing.assert_array_equal(res["eroded"][i], exp["eroded"])
            # the fast predicate path (no float plane requested) must agree with the one above
            res2 = ctx.process_batch(raws, nm03.default_params(**kw), want=("dilated",))
            np.testing.assert_array_equal(res2["dilated"], res["dilated"])
            assert res["rg_mask"].sum() > 0
    print("variant ok")
""") % (cm.ROOT, os.path.join(cm.ROOT, "tests"))

VARIANTS = {
    "plain_loads": {"NM03_NO_TMA": "1"},
    "no_skip_list": {"NM03_NO_SKIP": "1"},
    "no_tile_order": {"NM03_NO_TILE_ORDER": "1"},
    "static_work": {"NM03_NO_DYNAMIC": "1"},
    "exact_sharpen": {"NM03_FORCE_EXACT_SHARPEN": "1"},
    "cluster_of_8": {"NM03_CLUSTER": "8"},
    "cluster_of_1": {"NM03_CLUSTER": "1"},
    "everything_off": {"NM03_NO_TMA": "1", "NM03_NO_SKIP": "1", "NM03_NO_TILE_ORDER": "1", "NM03_NO_DYNAMIC": "1",
                       "NM03_FORCE_EXACT_SHARPEN": "1"},
}


@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_variant_matches_oracle(name):
    env = dict(os.environ, **VARIANTS[name])
    r = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "variant ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def test_zero_border_with_negative_values_is_refused_not_wrong(lib):
    """The zero-border mode carries out-of-image samples as the lowest key, which is only the place of 0.0 in a sorted
    window when clip(normalise()) is non-decreasing and never negative; other parameter sets are refused."""
    raw = cm.phantom(128, 128, 1, 1)
    with lib.Context([0]) as ctx:
        for kw in (dict(clip_min=-3.0, norm_low=-1.0), dict(norm_low=2.5, norm_high=0.5)):
            with pytest.raises(lib.Nm03Error) as e:
                ctx.preprocess(raw, lib.default_params(border_mode=1, **kw))
            assert e.value.code == lib.ERR_UNSUPPORTED and "zero border" in str(e.value)
        got = ctx.preprocess(raw, lib.default_params(border_mode=1, clip_min=0.0))      # 0.0 itself is fine
    f = cm.o_clip(cm.o_normalize(raw), 0.0, 4000.0)
    np.testing.assert_array_equal(got["median"], cm.o_median(f, 7, 1))


def test_device_entry_accepts_float_planes_at_any_4_byte_offset(lib):
    """A caller may hand in a view into a larger buffer: the float outputs then start on a 4-byte, not a 16-byte, boundary."""
    import torch
    raw = np.stack([cm.phantom(256, 256, 6, s) for s in range(2)])
    with lib.Context([0]) as ctx:
        d_in = torch.from_numpy(raw).cuda()
        n = raw.size
        flat_s = torch.zeros(n + 8, dtype=torch.float32, device="cuda")
        flat_m = torch.zeros(n + 8, dtype=torch.float32, device="cuda")
        d_mask = torch.zeros(raw.shape, dtype=torch.uint8, device="cuda")
        res = {}
        for off in (0, 1, 3):
            o = lib.Outputs()
            o.sharpened = flat_s.data_ptr() + 4 * off
            o.median = flat_m.data_ptr() + 4 * off
            o.dilated = d_mask.data_ptr()
            ctx.process_batch_device(d_in.data_ptr(), 2, 256, 256, lib.default_params(), o)
            torch.cuda.synchronize()
            res[off] = (flat_s[off:off + n].cpu().numpy().copy(), flat_m[off:off + n].cpu().numpy().copy())
    rc, exp = cm.o_pipeline(raw[0])
    np.testing.assert_array_equal(res[0][0][:256 * 256].reshape(256, 256), exp["sharpened"])
    for off in (1, 3):
        np.testing.assert_array_equal(res[off][0], res[0][0])
        np.testing.assert_array_equal(res[off][1], res[0][1])
