"""Slices beyond what one thread-block cluster can keep in shared memory take the global-memory
fallback of the flood fill (band-wise passes relaunched until a device flag stays clear).  Checked
(a) on a genuinely large slice and (b) by forcing the fallback for ordinary sizes in a fresh process
(the switch is read when the library is loaded)."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import scipy.ndimage as ndi

import common as cm

pytestmark = pytest.mark.gpu


def test_region_grow_and_morphology_on_a_slice_beyond_cluster_capacity(lib):
    h, w = 3100, 3300   # 104 words x 3100 rows x 2 planes = 2.6 MB > 8 CTAs x 200 KB
    rng = np.random.default_rng(5)
    img = ndi.zoom(ndi.gaussian_filter(rng.uniform(0, 1, (h // 20, w // 20)), 2.0), 20, order=1)[:h, :w].astype(np.float32)
    img = (img - img.min()) / (img.max() - img.min())
    seeds = cm.o_make_seeds(w, h)
    with lib.Context([0]) as ctx:
        for conn in (4, 8):
            got, it = ctx.region_grow(img, seeds, lib.default_params(rg_min=0.4, rg_max=0.6, rg_connectivity=conn))
            rc, exp, _ = cm.o_region_grow(img, 0.4, 0.6, seeds, conn=conn)
            np.testing.assert_array_equal(got, exp)
            assert it >= 2 and exp.sum() > 10000
        m = exp
        np.testing.assert_array_equal(ctx.morphology(m, 7, 0, 0), cm.o_dilate(m, 7)[1])
        np.testing.assert_array_equal(ctx.morphology(m, 5, 0, 1), cm.o_erode(m, 5)[1])


def test_forced_fallback_matches_oracle_in_a_fresh_process():
    code = textwrap.dedent("""
        import sys, numpy as np
        sys.path.insert(0, %r); sys.path.insert(0, %r)
        import common as cm
        nm03 = cm.load_product()
        with nm03.Context([0]) as ctx:
            raws = np.stack([cm.phantom(512, 512, 4, i, kind="lowcontrast" if i == 2 else "default") for i in range(3)])
            for kw in (dict(erode_size=3), dict(rg_connectivity=8, rg_border_expand=0, dilate_size=5), dict(rg_mode=1)):
                res = ctx.process_batch(raws, nm03.default_params(**kw), want=("dilated", "rg_mask", "eroded", "rg_iterations"))
                for i in range(3):
                    rc, exp = cm.o_pipeline(raws[i], cm.oparams(**kw))
                    assert rc == 0
                    np.testing.assert_array_equal(res["rg_mask"][i], exp["rg_mask"])
                    np.testing.assert_array_equal(res["dilated"][i], exp["dilated"])
                    if kw.get("erode_size"):
                        np.testing.assert_array_equal(res["eroded"][i], exp["eroded"])
                assert res["rg_iterations"].min() >= 2
        print("forced fallback ok")
    """) % (cm.ROOT, os.path.join(cm.ROOT, "tests"))
    env = dict(os.environ, NM03_FORCE_GLOBAL_FILL="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "forced fallback ok" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
