#!/usr/bin/env python3
"""Per-stage dumps of the reference pipeline from a real FAST install -> tests/golden/fast/*.npz.

NOT runnable in this repository's containers (no FAST, no OpenCL, no network): it is the recipe for a maintainer who
has `pyfast` (pip install pyfast, FAST >= 4.x).  The calls mirror the reference one to one
(src/sequential/main_sequential.cpp:175-252; erosion branch src/test/test_pipeline.cpp:119-121); the pyfast spelling
of each call is written from memory of FAST's Python wrappers and may need a touch-up on the first run.

  python tools/fast_dump.py --out tests/golden/fast                 # phantom slices 256x256 and 512x512
  python tools/fast_dump.py --out tests/golden/fast --dicom a.dcm   # or any single-frame 16-bit DICOM slices

What it pins once the files exist (tests/test_fast_pins.py): the sampler border mode of median and sharpen, the
multiply-add order of the unsharp mask, OpenCL vs host region growing, 4- vs 8-connectivity, whether image-border
pixels expand, and disc vs square structuring elements — every switch SURVEY §8c could not verify offline.
"""
import argparse
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def seeds_for(w, h):
    """main_sequential.cpp:214-241."""
    cx, cy, ox, oy = w // 2, h // 2, w // 8, h // 8
    pts = [(cx, cy), (cx + ox, cy), (cx - ox, cy), (cx, cy + oy), (cx, cy - oy)]
    for x in range(w // 4, w * 3 // 4, w // 10):
        for y in range(h // 4, h * 3 // 4, h // 10):
            pts.append((x, y))
    return pts


def dump_one(fast, path, out_path):
    importer = fast.DICOMFileImporter.create(path)
    importer.setLoadSeries(False)
    raw_img = importer.runAndGetOutputData()
    w, h = raw_img.getWidth(), raw_img.getHeight()
    normalize = fast.IntensityNormalization.create(0.5, 2.5, 0.0, 10000.0).connect(importer)      # :195-198
    clipping = fast.IntensityClipping.create(0.68, 4000.0).connect(normalize)                       # :200-202
    median = fast.VectorMedianFilter.create(7).connect(clipping)                                    # :204-206
    sharpen = fast.ImageSharpening.create(2.0, 0.5, 9).connect(median)                              # :208-210
    seeds = [fast.Vector3i(x, y, 0) if hasattr(fast, "Vector3i") else (x, y, 0) for x, y in seeds_for(w, h)]
    rg = fast.SeededRegionGrowing.create(0.74, 0.91, seeds).connect(sharpen)                        # :232-243
    caster = fast.ImageCaster.create(fast.TYPE_UINT8).connect(rg)                                   # :246-248
    dilation = fast.Dilation.create(3).connect(caster)                                              # :250-252
    erosion = fast.Erosion.create(3).connect(caster)                                                # test_pipeline.cpp:119-121
    grab = lambda po: np.asarray(po.runAndGetOutputData()).reshape(h, w)                            # noqa: E731
    np.savez_compressed(out_path, raw=grab(importer), median=grab(median).astype(np.float32),
                        sharpened=grab(sharpen).astype(np.float32), rg_mask=grab(rg).astype(np.uint8),
                        dilated=grab(dilation).astype(np.uint8), eroded=grab(erosion).astype(np.uint8),
                        fast_version=getattr(fast, "__version__", "unknown"))
    print("wrote", out_path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "fast"))
    ap.add_argument("--dicom", nargs="*", default=[])
    a = ap.parse_args()
    try:
        import fast
    except ImportError:
        sys.exit("pyfast is not installed: this script needs a real FAST install (see tests/golden/fast/README.md)")
    os.makedirs(a.out, exist_ok=True)
    files = list(a.dicom)
    if not files:
        drv = os.path.join(ROOT, "nm03-capstone-project_b200", "host", "nm03_driver")
        td = tempfile.mkdtemp()
        for size in (256, 512):
            subprocess.check_call([drv, "--make-dataset", f"{td}/{size}", "--patients", "2", "--slices", "3", "--size", str(size)])
            subprocess.check_call([drv, "--make-dataset", f"{td}/{size}lc", "--patients", "1", "--slices", "2", "--size", str(size), "--lowcontrast"])
        for root, _, names in os.walk(td):
            files += [os.path.join(root, n) for n in sorted(names) if n.endswith(".dcm")]
    for i, f in enumerate(files):
        dump_one(fast, f, os.path.join(a.out, f"fast_{i:03d}.npz"))


if __name__ == "__main__":
    main()
