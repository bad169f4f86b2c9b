"""The ticketed asynchronous host-buffer entry point (nm03_process_batch_async / nm03_wait, SURVEY §8b "async variant
with stream ticket") and the bit-packed result (nm03_outputs.dilated_bits), through the C ABI, against the blocking
entry point and the oracle."""
import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu


def _series(n, w, h, patient, kind="default"):
    return np.stack([cm.phantom(w, h, patient, s, kind=kind) for s in range(n)])


def test_one_thread_keeps_three_batches_in_flight(lib):
    with lib.Context([0]) as ctx:
        batches = [ctx.host_alloc((6, 256, 256), np.int16) for _ in range(3)]
        outs = [ctx.host_alloc((6, 256, 256), np.uint8) for _ in range(3)]
        for b, buf in enumerate(batches):
            buf[:] = _series(6, 256, 256, 40 + b, "lowcontrast" if b == 1 else "default")
        for _ in range(4):   # slots are reused round after round
            for o in outs:
                o[:] = 7
            tickets = [ctx.process_batch_async(batches[b], lib.default_params(), {"dilated": outs[b]}) for b in range(3)]
            for t in reversed(tickets):          # waiting in another order than queueing is allowed
                ctx.wait(t)
            for b in range(3):
                np.testing.assert_array_equal(outs[b], cm.o_batch_omp(np.array(batches[b])))
        with pytest.raises(ValueError):
            ctx.wait(tickets[0])                 # a ticket is waited for once


def test_async_matches_blocking_call_with_all_outputs(lib):
    raw = _series(5, 200, 150, 3)
    want = ("dilated", "eroded", "rg_mask", "sharpened", "median", "rg_iterations", "dilated_bits")
    with lib.Context([0]) as ctx:
        p = lib.default_params(erode_size=3)
        a = ctx.process_batch(raw, p, want=want)
        t = ctx.process_batch_async(raw, p, want=want)
        b = ctx.wait(t)
    for name in want:
        np.testing.assert_array_equal(a[name], b[name], err_msg=name)
    np.testing.assert_array_equal(lib.unpack_bits(a["dilated_bits"], 200), a["dilated"])


def test_more_tickets_than_slots_is_refused_not_blocked(lib):
    raw = _series(4, 128, 128, 9)
    with lib.Context([0]) as ctx:
        tickets = []
        with pytest.raises(lib.Nm03Error) as ei:
            for _ in range(9):
                tickets.append(ctx.process_batch_async(raw, lib.default_params()))
        assert ei.value.code == lib.ERR_BUSY and len(tickets) == 8
        exp = cm.o_batch_omp(raw)
        for t in tickets:
            np.testing.assert_array_equal(ctx.wait(t)["dilated"], exp)
        ctx.wait(ctx.process_batch_async(raw, lib.default_params()))   # slots are free again


def test_async_errors_come_back_without_a_ticket(lib):
    with lib.Context([0]) as ctx:
        with pytest.raises(lib.Nm03Error) as ei:
            ctx.process_batch_async(np.zeros((2, 96, 96), np.int16), lib.default_params())
        assert ei.value.code == lib.ERR_TOO_SMALL and "Image dimensions too small: 96x96" in str(ei.value)
        with pytest.raises(lib.Nm03Error) as ei:
            ctx.process_batch_async(np.zeros((2, 128, 128), np.int16), lib.default_params(dilate_size=4))
        assert ei.value.code == lib.ERR_PARAM
        # the slot of the failed call was handed back: eight more batches can still be queued
        ts = [ctx.process_batch_async(_series(2, 128, 128, 1), lib.default_params()) for _ in range(8)]
        for t in ts:
            ctx.wait(t)


@pytest.mark.parametrize("w,h,size", [(512, 512, 3), (130, 101, 3), (256, 160, 7), (100, 100, 15)])
def test_bit_packed_result_equals_the_byte_mask(lib, w, h, size):
    raw = _series(3, w, h, 11)
    with lib.Context([0]) as ctx:
        got = ctx.process_batch(raw, lib.default_params(dilate_size=size), want=("dilated", "dilated_bits"))
        only = ctx.process_batch(raw, lib.default_params(dilate_size=size), want=("dilated_bits",))
    exp = cm.o_batch_omp(raw, cm.oparams(dilate_size=size))
    np.testing.assert_array_equal(got["dilated"], exp)
    np.testing.assert_array_equal(lib.unpack_bits(got["dilated_bits"], w), exp)
    np.testing.assert_array_equal(only["dilated_bits"], got["dilated_bits"])
    if w % 32:   # bits beyond the last pixel of a row are zero
        assert (got["dilated_bits"][..., -1] >> (w % 32)).max() == 0


def test_bit_packed_result_from_the_global_memory_fallback(lib, monkeypatch):
    import subprocess, sys, os   # the fallback is selected by an environment switch read at load time: use a child process
    code = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r); import numpy as np, common as cm;"
            "lib = cm.load_product(); raw = np.stack([cm.phantom(256, 256, 2, s) for s in range(2)]);"
            "ctx = lib.Context([0]); g = ctx.process_batch(raw, lib.default_params(), want=('dilated', 'dilated_bits'));"
            "assert (lib.unpack_bits(g['dilated_bits'], 256) == g['dilated']).all() and g['dilated'].sum() > 0;"
            "assert (g['dilated'] == cm.o_batch_omp(raw)).all(); print('ok')") % (cm.ROOT, os.path.join(cm.ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, NM03_FORCE_GLOBAL_FILL="1"), capture_output=True, text=True)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]


def test_status_is_the_result_of_the_call_for_every_slice(lib):
    with lib.Context([0]) as ctx:
        ok = ctx.process_batch(_series(3, 128, 128, 5), lib.default_params())
        assert (ok["status"] == 0).all()
        res = {}
        with pytest.raises(lib.Nm03Error):
            ctx.process_batch(np.zeros((3, 64, 64), np.int16), lib.default_params(), out=res)
        assert (res["status"] == lib.ERR_TOO_SMALL).all()


def test_params_object_is_not_modified_by_the_wrapper(lib):
    p = lib.default_params()
    with lib.Context([0]) as ctx:
        u = cm.phantom(128, 128, 0, 0, dtype=np.uint16)
        ctx.process_batch(u[None], p)
        assert p.input_unsigned == 0
        s = ctx.process_batch(u.astype(np.int16)[None], p)["dilated"]
    np.testing.assert_array_equal(s[0], cm.o_pipeline(u.astype(np.int16))[1]["dilated"])


def test_work_counter_counts_sorted_median_items(lib):
    import torch
    raw = _series(4, 512, 512, 13)
    with lib.Context([0]) as ctx:
        d_in = torch.from_numpy(raw).cuda()
        d_out = torch.empty((4, 512, 512), dtype=torch.uint8, device="cuda")
        o = lib.Outputs()
        o.dilated = d_out.data_ptr()
        ctx.process_batch_device(d_in.data_ptr(), 4, 512, 512, lib.default_params(), o)
        ctx.profile_work()
        ctx.process_batch_device(d_in.data_ptr(), 4, 512, 512, lib.default_params(), o)
        n = ctx.profile_work()
        assert ctx.profile_work() == 0
    all_items = 4 * 32 * 36 * 17            # slices x tiles x packed rows x runs of 8 columns
    assert 0.3 * all_items < n < all_items, (n, all_items)   # the air around the head takes the clip-floor shortcut
    assert 1000 < lib.median_network_ops(7) < 1500 and lib.median_network_ops(4) == 0


def test_prepared_parameter_blocks_are_keyed_by_everything_they_depend_on(lib):
    """The library caches the prepared kernel parameter blocks per (parameters, size, seeds, requested planes).  One context,
    parameter sets and sizes interleaved over several rounds — every call must give what a fresh context gives: the oracle's."""
    rng = np.random.default_rng(3)
    cases = [
        (_series(2, 256, 256, 21), dict(), None, ("dilated",)),
        (_series(2, 256, 256, 21), dict(median_window=5, dilate_size=7), None, ("dilated",)),
        (_series(2, 200, 150, 22), dict(), None, ("dilated", "sharpened")),
        (_series(2, 256, 256, 21), dict(), np.array([[128, 128], [100, 90], [5, 5]], np.int32), ("dilated",)),
        (_series(2, 256, 256, 21), dict(), np.array([[5, 5], [10, 10], [6, 5]], np.int32), ("dilated",)),   # same count, other points: empty mask
        (_series(2, 256, 256, 23, "lowcontrast"), dict(erode_size=3), None, ("dilated", "eroded")),
        (_series(2, 256, 256, 21).astype(np.uint16), dict(), None, ("dilated",)),
    ]
    expected = []
    for raw, kw, seeds, want in cases:
        p = cm.oparams(**kw)
        if raw.dtype == np.uint16:
            p.input_unsigned = 1
        exp = []
        for i in range(len(raw)):
            if seeds is None:
                rc, one = cm.o_pipeline(raw[i], p)
            else:
                f = cm.o_clip(cm.o_normalize(raw[i]))
                sh = cm.o_sharpen(cm.o_median(f, 7))
                rc, rg, _ = cm.o_region_grow(sh, 0.74, 0.91, seeds)
                one = {"dilated": cm.o_dilate(rg, 3)[1], "sharpened": sh}
            assert rc == 0
            exp.append(one)
        expected.append(exp)
    with lib.Context([0]) as ctx:
        for rnd in range(4):
            for idx in rng.permutation(len(cases)):
                raw, kw, seeds, want = cases[idx]
                got = ctx.process_batch(raw, lib.default_params(**kw), seeds=seeds, want=want)
                for i in range(len(raw)):
                    for name in want:
                        np.testing.assert_array_equal(got[name][i], expected[idx][i][name], err_msg=f"round {rnd} case {idx} {name}")
