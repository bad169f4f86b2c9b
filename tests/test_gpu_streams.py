"""The device entry point called from several caller-owned streams at once: the library gives each calling
stream a scratch slot of its own (predicate plane, K1 work counters, LUT), so batches in flight on different
streams must not disturb each other, and a slot handed from one stream to another is ordered by its event."""
import numpy as np
import pytest

import common as cm

pytestmark = pytest.mark.gpu

W = H = 512


@pytest.mark.parametrize("n_streams", [2, 3, 10])   # 10 > the library's 8 slots: slots change hands between streams
def test_interleaved_streams_match_oracle_and_single_stream(lib, n_streams):
    import torch
    dev = torch.device("cuda", 0)
    rounds = 5
    with lib.Context([0]) as ctx:
        params = lib.default_params()
        batches = [np.stack([cm.phantom(W, H, 20 + b, i, kind="lowcontrast" if (b + i) % 3 == 0 else "default")
                             for i in range(6 + (b % 3))]) for b in range(n_streams)]
        d_in = [torch.from_numpy(x).to(dev) for x in batches]
        d_out = [torch.zeros((x.shape[0], H, W), dtype=torch.uint8, device=dev) for x in batches]
        outs = []
        for o in d_out:
            s = lib.Outputs()
            s.dilated = o.data_ptr()
            outs.append(s)
        # reference result: one batch at a time on one stream
        single = []
        st0 = torch.cuda.Stream(device=dev)
        for b in range(n_streams):
            ctx.process_batch_device(d_in[b].data_ptr(), batches[b].shape[0], W, H, params, outs[b], st0.cuda_stream)
            st0.synchronize()
            single.append(d_out[b].cpu().numpy())
            d_out[b].zero_()
        torch.cuda.synchronize()
        streams = [torch.cuda.Stream(device=dev) for _ in range(n_streams)]
        for _ in range(rounds):
            for b in range(n_streams):
                ctx.process_batch_device(d_in[b].data_ptr(), batches[b].shape[0], W, H, params, outs[b],
                                         streams[b].cuda_stream)
        torch.cuda.synchronize()
        for b in range(n_streams):
            np.testing.assert_array_equal(d_out[b].cpu().numpy(), single[b], err_msg=f"stream {b}")
        rc, exp = cm.o_pipeline(batches[1][2], cm.oparams())
        assert rc == 0
        np.testing.assert_array_equal(single[1][2], exp["dilated"])
        assert single[1][2].sum() > 0


def test_host_entry_after_device_entry_on_a_foreign_stream(lib):
    """A slot last used from a caller's stream is picked up by the host-buffer entry point on the slot's own stream."""
    import torch
    dev = torch.device("cuda", 0)
    with lib.Context([0]) as ctx:
        params = lib.default_params()
        raws = np.stack([cm.phantom(W, H, 31, i) for i in range(8)])
        d_in = torch.from_numpy(raws).to(dev)
        d_out = torch.zeros((8, H, W), dtype=torch.uint8, device=dev)
        o = lib.Outputs()
        o.dilated = d_out.data_ptr()
        st = torch.cuda.Stream(device=dev)
        for _ in range(4):
            ctx.process_batch_device(d_in.data_ptr(), 8, W, H, params, o, st.cuda_stream)
        res = ctx.process_batch(raws[::-1].copy(), params, want=("dilated",))   # queued while `st` is still busy
        torch.cuda.synchronize()
        np.testing.assert_array_equal(res["dilated"][::-1], d_out.cpu().numpy())


SPLIT_CODE = """
import sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
import common as cm
lib = cm.load_product()
W = H = 512
n = int(sys.argv[1])
dev = torch.device("cuda", 0)
a = np.stack([cm.phantom(W, H, 50, i, kind="lowcontrast" if i %% 5 == 0 else "default") for i in range(n)])
b = np.stack([cm.phantom(W, H, 51, i) for i in range(n)])
want = ("dilated", "eroded", "rg_mask", "dilated_bits", "rg_iterations")
with lib.Context([0]) as ctx:
    params = lib.default_params(erode_size=3)
    d_a, d_b = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
    bufs = {k: torch.zeros((n, H, W), dtype=torch.uint8, device=dev) for k in ("dilated", "eroded", "rg_mask")}
    bufs["dilated_bits"] = torch.zeros((n, H, W // 32), dtype=torch.int32, device=dev)
    bufs["rg_iterations"] = torch.zeros(n, dtype=torch.int32, device=dev)
    o = lib.Outputs()
    for k in want:
        setattr(o, k, bufs[k].data_ptr())
    st = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(st):
        ctx.process_batch_device(d_b.data_ptr(), n, W, H, params, o, st.cuda_stream)   # first use of the slots builds their tables
        st.synchronize()
        launches = ctx.kernel_launches()
        ctx.process_batch_device(d_a.data_ptr(), n, W, H, params, o, st.cuda_stream)
        parts = (ctx.kernel_launches() - launches) // 2
        after_a = {k: bufs[k].clone() for k in want}          # queued on st behind call A
        ctx.process_batch_device(d_b.data_ptr(), n, W, H, params, o, st.cuda_stream)
    st.synchronize()
    assert parts == (4 if n >= 24 else 2), parts
    for raw, got in ((a, after_a), (b, bufs)):
        exp = cm.o_batch_omp(raw, cm.oparams(erode_size=3))
        np.testing.assert_array_equal(got["dilated"].cpu().numpy(), exp)
        np.testing.assert_array_equal(lib.unpack_bits(got["dilated_bits"].cpu().numpy().view(np.uint32), W), exp)
        for i in (0, n - 1):
            rc, one = cm.o_pipeline(raw[i], cm.oparams(erode_size=3))
            np.testing.assert_array_equal(got["eroded"][i].cpu().numpy(), one["eroded"])
            np.testing.assert_array_equal(got["rg_mask"][i].cpu().numpy(), one["rg_mask"])
        assert int(got["rg_iterations"].min()) >= 1
print("ok")
"""


@pytest.mark.parametrize("n", [24, 13])
def test_opt_in_split_of_a_one_stream_batch_keeps_stream_order(n):
    """NM03_SPLIT=1 (off by default: measured slower): a lone single-stream caller's batch is cut into parts that run on
    library streams, fork / join on the caller's stream.  Results must be the oracle's, and the caller's stream must still
    order everything: a copy queued behind call A sees A's masks, and call B may then overwrite the same output buffer."""
    import os, subprocess, sys
    code = SPLIT_CODE % (cm.ROOT, os.path.join(cm.ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", code, str(n)], env=dict(os.environ, NM03_SPLIT="1"), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-3000:]
