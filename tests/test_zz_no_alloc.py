"""The allocation-free pipeline entry point, nm03_process_batch_device_noalloc.

Some callers of this library cannot allocate device memory while they run: feeder threads whose
cudaMalloc would stall every other stream of the process, and processes whose GPU memory is already
owned by someone else's pool.  For them the library keeps the scratch bit-plane between its two
kernels in a fixed 1 MiB array of its own.  This file checks, through the C ABI, that the variant
gives exactly the masks of the regular device entry point (masks are compared bit for bit, as in the
other GPU tests) on two calls in a row — the second call finds the first call's bits still in the
shared array — and that it refuses batches that do not fit.

Named test_zz_* so that pytest collects it last: a stop-at-first-failure run has then already run
every other test.  It creates its own context and tensors, so it passes alone as well as after the
rest of the suite.
"""
import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu

W = H = 512


def _run(nm03, ctx, torch, fn, d_in, params, want_erode):
    n = d_in.shape[0]
    dev = d_in.device
    o = nm03.Outputs()
    dil = torch.zeros((n, H, W), dtype=torch.uint8, device=dev)
    rg = torch.zeros((n, H, W), dtype=torch.uint8, device=dev)
    ero = torch.zeros((n, H, W), dtype=torch.uint8, device=dev)
    o.dilated, o.rg_mask = dil.data_ptr(), rg.data_ptr()
    if want_erode:
        o.eroded = ero.data_ptr()
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        fn(d_in.data_ptr(), n, W, H, params, o, stream.cuda_stream)
    stream.synchronize()
    return dil.cpu().numpy(), rg.cpu().numpy(), ero.cpu().numpy()


def test_noalloc_matches_regular_entry_point_twice():
    import torch
    nm03 = cm.load_product()
    dev = torch.device("cuda", 0)
    with nm03.Context([0]) as ctx:
        params = nm03.default_params(erode_size=3)
        # two different batches, back to back: the second call reuses the array the first one filled
        first = np.stack([cm.phantom(W, H, 2, i) for i in range(5)])
        second = np.stack([cm.phantom(W, H, 9, i, kind="lowcontrast" if i % 2 else "default") for i in range(3)])
        for raws in (first, second):
            d_in = torch.from_numpy(raws).to(dev)
            torch.cuda.synchronize()
            got = _run(nm03, ctx, torch, ctx.process_batch_device_noalloc, d_in, params, True)
            exp = _run(nm03, ctx, torch, ctx.process_batch_device, d_in, params, True)
            for g, e, name in zip(got, exp, ("dilated", "rg_mask", "eroded")):
                np.testing.assert_array_equal(g, e, err_msg=name)
            assert got[0].sum() > 0
        # and against the CPU oracle for one slice of the second batch
        rc, o = cm.o_pipeline(second[1], cm.oparams(erode_size=3))
        assert rc == 0
        np.testing.assert_array_equal(got[0][1], o["dilated"])


def test_noalloc_rejects_batches_beyond_the_builtin_scratch():
    import torch
    nm03 = cm.load_product()
    dev = torch.device("cuda", 0)
    per_slice = H * ((W + 31) // 32)
    too_many = nm03.NOALLOC_MAX_PRED_WORDS // per_slice + 1
    with nm03.Context([0]) as ctx:
        d_in = torch.zeros((too_many, H, W), dtype=torch.int16, device=dev)
        dil = torch.zeros((too_many, H, W), dtype=torch.uint8, device=dev)
        o = nm03.Outputs()
        o.dilated = dil.data_ptr()
        with pytest.raises(nm03.Nm03Error) as e:
            ctx.process_batch_device_noalloc(d_in.data_ptr(), too_many, W, H, nm03.default_params(), o, 0)
        assert e.value.code == nm03.ERR_TOO_LARGE
        assert "at most 32 slices" in str(e.value)
        # the largest batch that fits is accepted
        ctx.process_batch_device_noalloc(d_in.data_ptr(), too_many - 1, W, H, nm03.default_params(), o, 0)
        torch.cuda.synchronize()
