"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/nm03.h declares,
and its host-only entry points behave (no GPU compute here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import common as cm


def test_exports_match_header(lib):
    hdr = open(os.path.join(cm.ROOT, "include", "nm03.h")).read()
    declared = set(re.findall(r"\b(nm03_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(lib.EXPORTS), declared ^ set(lib.EXPORTS)
    so = lib.lib()
    for name in declared:
        assert hasattr(so, name), name


def test_default_params_are_reference_literals(lib):
    p = lib.default_params()
    # main_sequential.cpp:196,200,204,208,233,250
    assert (p.norm_low, p.norm_high, p.norm_min, p.norm_max) == (0.5, 2.5, 0.0, 10000.0)
    assert p.clip_min == pytest.approx(0.68) and p.clip_max == 4000.0
    assert p.median_window == 7 and p.sharpen_mask == 9 and p.sharpen_gain == 2.0 and p.sharpen_sigma == 0.5
    assert p.rg_min == pytest.approx(0.74) and p.rg_max == pytest.approx(0.91)
    assert p.dilate_size == 3 and p.erode_size == 0 and p.rg_connectivity == 4
    # the oracle's parameter block has the same layout and defaults
    o = cm.oparams()
    assert bytes(p) == bytes(o)


@pytest.mark.parametrize("w,h", [(100, 100), (256, 256), (512, 512), (1024, 1024), (320, 200)])
def test_seeds_match_oracle(lib, w, h):
    np.testing.assert_array_equal(lib.make_seeds(w, h), cm.o_make_seeds(w, h))


def test_seeds_too_small(lib):
    with pytest.raises(lib.Nm03Error) as e:
        lib.make_seeds(9, 500)
    assert e.value.code == lib.ERR_TOO_SMALL


def test_init_without_gpu_fails_loudly(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lib.Nm03Error) as e:
        lib.Context([0])
    assert e.value.code == lib.ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)
