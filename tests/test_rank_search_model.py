"""The selection rule of the generic median paths (K1's rank search for windows 11-15, `k_median_big` above), restated in
numpy: most-significant-bit first, starting below the highest bit in which the window's smallest and largest keys differ, a
trial value t is accepted when fewer than (K*K+1)/2 samples lie below it.  Checked against a sort, ties and all 16 bits
included.  (That the kernels implement this rule is what the GPU parity tests check against the oracle.)"""
import numpy as np
import pytest


def rank_search_median(keys):
    """keys: unsigned-order 16-bit keys of one window."""
    rank = (len(keys) + 1) // 2
    mn, mx = int(keys.min()), int(keys.max())
    cand = mn
    if mn != mx:
        hb = (mn ^ mx).bit_length() - 1
        cand = mn & ~((2 << hb) - 1)
        for b in range(hb, -1, -1):
            t = cand | (1 << b)
            if int((keys < t).sum()) < rank:
                cand = t
    return cand


@pytest.mark.parametrize("k", [3, 11, 15, 17, 33])
def test_rank_search_equals_sorted_median(k):
    rng = np.random.default_rng(k)
    for _ in range(400):
        span = int(rng.choice([1, 2, 5, 100, 4096, 65536]))
        lo = int(rng.integers(0, 65536 - span + 1))
        keys = rng.integers(lo, lo + span, k * k)
        assert rank_search_median(keys) == int(np.sort(keys)[(k * k) // 2])


def test_signed_samples_through_the_key_flip():
    rng = np.random.default_rng(1)
    for _ in range(200):
        v = rng.integers(-32768, 32768, 49).astype(np.int16)
        keys = v.astype(np.int32) + 32768            # signed sample ^ 0x8000 = its rank among unsigned keys
        assert rank_search_median(keys) - 32768 == int(np.sort(v)[24])
