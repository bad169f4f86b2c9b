"""Committed fixtures (tests/golden/*.npz, produced by tests/golden/make_golden.py from the oracle).
CPU: the oracle still reproduces them.  GPU: the CUDA path reproduces them without the oracle."""
import glob
import os

import numpy as np
import pytest

import common as cm

FILES = sorted(glob.glob(os.path.join(cm.ROOT, "tests", "golden", "*.npz")))


def _load(path):
    z = np.load(path, allow_pickle=True)
    kw = {k: int(v) for k, v in z["params"]} if z["params"].size else {}
    return z, kw


def _unpack(z, name, w):
    return np.unpackbits(z[name], axis=1)[:, :w]


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_oracle_reproduces_golden(path):
    z, kw = _load(path)
    raw = z["raw"]
    rc, r = cm.o_pipeline(raw, cm.oparams(**kw))
    assert rc == 0
    np.testing.assert_array_equal(r["median"].view(np.uint32), z["median_u32"])
    np.testing.assert_array_equal(r["sharpened"].view(np.uint32), z["sharpened_u32"])
    for k in ("rg_mask", "dilated", "eroded"):
        if k in z.files:
            np.testing.assert_array_equal(r[k], _unpack(z, k, raw.shape[1]))


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_cuda_reproduces_golden(lib, path):
    z, kw = _load(path)
    raw = z["raw"]
    want = ("dilated", "rg_mask", "sharpened", "median") + (("eroded",) if "eroded" in z.files else ())
    with lib.Context([0]) as ctx:
        res = ctx.process_batch(raw[None], lib.default_params(**kw), want=want)
    np.testing.assert_array_equal(res["median"][0].view(np.uint32), z["median_u32"])
    np.testing.assert_array_equal(res["sharpened"][0].view(np.uint32), z["sharpened_u32"])
    for k in ("rg_mask", "dilated", "eroded"):
        if k in z.files:
            np.testing.assert_array_equal(res[k][0], _unpack(z, k, raw.shape[1]))
