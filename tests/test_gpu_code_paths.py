"""The library's alternative code paths, each forced by its environment switch in a fresh process (the switches are
read when the library is loaded): plain loads instead of TMA, no clip-floor skip list, no central-tiles-first
order, static instead of counter-based work distribution, the exact sharpen phase instead of the fast one, a
fixed cluster size, the warp-specialised K1, and all of them together.  Every variant must give the oracle's masks bit for bit."""
import os
import subprocess
import sys
import textwrap

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
    "warp_specialised_k1": {"NM03_K1_WS": "1"},
    "everything_off": {"NM03_NO_TMA": "1", "NM03_NO_SKIP": "1", "NM03_NO_TILE_ORDER": "1", "NM03_NO_DYNAMIC": "1",
                       "NM03_FORCE_EXACT_SHARPEN": "1"},
}


@pytest.mark.parametrize("name", sorted(VARIANTS))
def test_variant_matches_oracle(name):
    env = dict(os.environ, **VARIANTS[name])
    r = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "variant ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
