"""Pins the oracle to FAST — once somebody with a FAST install has run tools/fast_dump.py.

For every tests/golden/fast/*.npz (per-stage outputs of the real reference pipeline) the oracle's switch space is
searched for the combinations that reproduce the dump: median and masks bit-exact, the sharpened plane within the
stated tolerance (1e-5 of the normalised range).  The test fails if the repository's DEFAULT switches are not among
them, and names a combination that is.  With no dumps (this repository's containers cannot run FAST) it is skipped
and parity stays "unpinned"."""
import glob
import itertools
import os

import numpy as np
import pytest

import common as cm

DUMPS = sorted(glob.glob(os.path.join(cm.ROOT, "tests", "golden", "fast", "*.npz")))
SHARPEN_TOL = 1e-5 * (2.5 - 0.5)
SPACE = dict(border_mode=(0, 1), sharpen_fma=(0, 1), rg_mode=(0, 1), rg_connectivity=(4, 8), rg_border_expand=(1, 0),
             se_shape=(0, 1))


def matching_switches(d):
    raw = d["raw"]
    hits = []
    for combo in itertools.product(*SPACE.values()):
        kw = dict(zip(SPACE, combo))
        rc, got = cm.o_pipeline(raw, cm.oparams(erode_size=3, **kw))
        if rc != 0:
            continue
        if (np.array_equal(got["median"], d["median"]) and np.max(np.abs(got["sharpened"] - d["sharpened"])) <= SHARPEN_TOL
                and np.array_equal(got["rg_mask"], d["rg_mask"]) and np.array_equal(got["dilated"], d["dilated"])
                and ("eroded" not in d or np.array_equal(got["eroded"], d["eroded"]))):
            hits.append(kw)
    return hits


@pytest.mark.skipif(not DUMPS, reason="no FAST dumps under tests/golden/fast/ (parity unpinned; see its README.md)")
@pytest.mark.parametrize("path", DUMPS)
def test_default_switches_reproduce_fast(path):
    d = dict(np.load(path))
    hits = matching_switches(d)
    assert hits, f"no switch combination of the oracle reproduces {os.path.basename(path)}: the restatement itself is off"
    default = {k: v[0] for k, v in SPACE.items()}
    assert default in hits, f"FAST behaves like {hits[0]}, the repository's defaults are {default}"


def test_switch_search_finds_the_oracle_itself():
    """Self-check of the search (runs everywhere): a dump made by the oracle with non-default switches is recognised."""
    raw = cm.phantom(128, 128, 3, 1)
    kw = dict(border_mode=1, sharpen_fma=1, rg_mode=0, rg_connectivity=8, rg_border_expand=1, se_shape=1)
    rc, d = cm.o_pipeline(raw, cm.oparams(erode_size=3, **kw))
    assert rc == 0
    d["raw"] = raw
    assert kw in matching_switches(d)
