"""The error bound of K1's fast sharpen predicate, checked numerically on the CPU.

`fill_fast` (nm03-capstone-project_b200/csrc/nm03_api.cu) promises that the kernel's separable FMA evaluation of the
sharpened value differs from the oracle's tap-by-tap value by at most eps = ea*|blur_a| + eb*|v| + ec, so that a pixel
farther than eps from both region-growing thresholds has the oracle's predicate bit (the others are recomputed exactly in
the kernel).  This test restates the bound and the kernel's arithmetic in numpy — fp32 values, every FMA rounded once — and
compares with the oracle's sharpen on images, parameters and both window choices (5 taps / all taps).  It checks the
mathematics of the bound; that the kernel computes what is restated here is what the GPU parity tests check."""
import numpy as np
import pytest

import common as cm

f32 = np.float32
U = 1.0 / 16777216.0


def fma(a, b, c):
    """One rounding: exact in float64 for fp32 operands of these magnitudes up to far below the bound's resolution."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


def taps(sigma, m):
    hs = (m - 1) // 2
    g = np.array([np.exp(f32(-float(x * x)) / f32(f32(2.0) * f32(sigma) * f32(sigma))) for x in range(-hs, hs + 1)]).astype(np.float32)
    gs = f32(0)
    for v in g:
        gs = f32(gs + v)
    return (g / gs).astype(np.float32)


def bound(gain, sigma, m, he, vmax, nonneg, rlo, rhi):
    """ea, eb, ec as fill_fast computes them (doubles, inflated the same way)."""
    hs = (m - 1) // 2
    mask = cm.o_gauss_mask(sigma, m).astype(np.float64)
    gx = taps(sigma, m)
    gamma = lambda n: n * U / (1 - n * U)  # noqa: E731
    rho = d_abs = d_out = 0.0
    for x in range(-hs, hs + 1):
        for y in range(-hs, hs + 1):
            wxy = float(mask[y + hs][x + hs])
            if abs(x) <= he and abs(y) <= he:
                sep = float(gx[x + hs]) * float(gx[y + hs])
                if wxy >= 1e-12:
                    rho = max(rho, abs(sep - wxy) / wxy)
                else:
                    d_abs += abs(sep - wxy)
            else:
                d_out += wxy
    gM, g2 = gamma(m * m), gamma(2.0 * (2 * he + 1))
    c1 = gM + g2 * (1 + rho) + rho + 2 * U
    G = abs(gain)
    ec = G * ((1 + gM) * d_out + d_abs) * vmax * 1.05
    if nonneg:
        c1 *= (1 + 2 * rho + 4 * g2) * 1.05
        ea = G * c1 + 6.3 * U * G
    else:
        ec += G * c1 * vmax * 1.05
        ea = 6.3 * U * G
    eb = U * (6.3 * G + 2.1)
    ea, eb, ec = ea + 4.2 * U * G, eb + 4.2 * U * (1 + G), ec + 4.2 * U * (abs(rlo) + abs(rhi))
    return gx, ea * 1.02, eb * 1.02, ec * 1.02


def fast_sharpen(med, gx, hs, he, gain):
    """Phase C of K1 in fast mode: per row a chain of FMAs over the taps, then a chain over the rows with the taps scaled
    by -gain, then one FMA with fl(1 + gain)."""
    h, w = med.shape
    pad = np.pad(med, hs, mode="edge")
    hsum = np.zeros((h + 2 * hs, w), np.float32)
    for t in range(-he, he + 1):
        hsum = fma(gx[t + hs], pad[:, hs + t:hs + t + w], hsum)
    gyn = (f32(-gain) * gx).astype(np.float32)
    nb = np.zeros((h, w), np.float32)
    for t in range(-he, he + 1):
        nb = fma(gyn[t + hs], hsum[hs + t:hs + t + h, :], nb)
    return fma(f32(1.0) + f32(gain), med, nb), nb


IMAGES = {
    "phantom": lambda: cm.phantom(192, 160, 3, 5),
    "lowcontrast": lambda: cm.phantom(160, 160, 4, 2, kind="lowcontrast"),
    "all16bits": lambda: np.random.default_rng(0).integers(-32768, 32768, (120, 130)).astype(np.int16),
    "ties": lambda: np.random.default_rng(1).integers(1195, 1206, (100, 90)).astype(np.int16),
    "steps": lambda: np.repeat(np.repeat(np.random.default_rng(7).integers(1190, 1211, (8, 9)), 16, 0), 16, 1).astype(np.int16),
}


@pytest.mark.parametrize("name", sorted(IMAGES))
@pytest.mark.parametrize("gain,sigma,m", [(2.0, 0.5, 9), (2.0, 0.8, 7), (-1.5, 0.8, 7), (2.0, 1.0, 9), (0.7, 0.4, 3), (3.0, 0.6, 5)])
def test_fast_value_stays_inside_the_band(name, gain, sigma, m):
    raw = IMAGES[name]()
    hs = (m - 1) // 2
    med = cm.o_median(cm.o_clip(cm.o_normalize(raw)), 7)
    exact = cm.o_sharpen(med, gain, sigma, m)
    ends = cm.o_clip(cm.o_normalize(np.array([[-32768, 32767]], np.int16)))
    vmax, nonneg = float(np.abs(ends).max()), bool((ends >= 0).all())
    for he in sorted({min(2, hs), hs}):
        gx, ea, eb, ec = bound(gain, sigma, m, he, vmax, nonneg, 0.74, 0.91)
        sv, nb = fast_sharpen(med, gx, hs, he, gain)
        eps = ea * (np.abs(nb.astype(np.float64)) / abs(gain)) + eb * np.abs(med.astype(np.float64)) + ec
        err = np.abs(sv.astype(np.float64) - exact.astype(np.float64))
        assert (err <= eps).all(), (he, float((err / eps).max()))
        if he == hs:   # with every tap the band is a few ulps wide: the fast path decides all but a sliver of pixels
            assert eps.max() < 3e-5 * max(1.0, float(np.abs(med).max()))
