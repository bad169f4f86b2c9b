#!/usr/bin/env python3
"""Regenerates tests/golden/*.npz from the CPU oracle (oracle/libnm03_oracle.so).

The reference ships no golden vectors, and FAST cannot run here, so these fixtures pin
the ORACLE (against accidental changes) and give the GPU tests inputs/outputs that travel to the
GPU box.  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import common as cm  # noqa: E402

CASES = {
    # name: (w, h, patient, slice, kind, oracle parameter overrides)
    "c1_test_path_256": (256, 256, 17, 14, "default", dict(erode_size=3)),
    "ragged_203x131_conn8": (203, 131, 3, 2, "default", dict(rg_connectivity=8, erode_size=5, dilate_size=5)),
    "lowcontrast_160": (160, 160, 5, 0, "lowcontrast", dict()),
    "median5_square_se": (192, 128, 7, 4, "default", dict(median_window=5, dilate_size=7, se_shape=1)),
}


def main():
    for name, (w, h, pat, sl, kind, kw) in CASES.items():
        raw = cm.phantom(w, h, pat, sl, kind)
        rc, r = cm.o_pipeline(raw, cm.oparams(**kw))
        assert rc == 0
        out = {"raw": raw, "params": np.array(sorted(kw.items()), dtype=object) if kw else np.array([], dtype=object)}
        out["median_u32"] = r["median"].view(np.uint32)          # exact bit patterns
        out["sharpened_u32"] = r["sharpened"].view(np.uint32)
        for k in ("rg_mask", "dilated", "eroded"):
            if k in r:
                out[k] = np.packbits(r[k], axis=1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, {k: int(r[k].sum()) for k in ("rg_mask", "dilated")})


if __name__ == "__main__":
    main()
