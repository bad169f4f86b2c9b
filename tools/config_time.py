#!/usr/bin/env python3
"""Device-resident throughput of the other BASELINE.json configurations on ONE GPU (their multi-GPU form shards
slices over ranks without any exchange, so the per-GPU figure scales as in profiles/r1g_scale.json):

  C4  large matrix: 1024x1024 slices, median window 5, dilation size 7 (radius 3)
  C5  worst-case region growing: 512x512 low-contrast phantom (grown region > 50 % of the slice)

Batches are issued round-robin on 4 streams as in bench.py; inputs rotate over more data than L2 holds.
usage: config_time.py [steps]"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common as cm  # noqa: E402

nm03 = importlib.import_module("nm03-capstone-project_b200")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
dev = torch.device("cuda", 0)


def run(name, w, h, n, kind, series, **kw):
    host = np.stack([np.stack([cm.phantom(w, h, p, s, kind=kind) for s in range(n)]) for p in range(series)])
    d_in = torch.from_numpy(host).to(dev)
    params = nm03.default_params(**kw)
    with nm03.Context([0]) as ctx:
        streams = [torch.cuda.Stream(device=dev) for _ in range(4)]
        d_out = torch.empty((4, n, h, w), dtype=torch.uint8, device=dev)
        d_it = torch.zeros(n, dtype=torch.int32, device=dev)
        outs = []
        for i in range(4):
            o = nm03.Outputs()
            o.dilated = d_out[i].data_ptr()
            outs.append(o)
        o_it = nm03.Outputs()
        o_it.dilated = d_out[0].data_ptr()
        o_it.rg_iterations = d_it.data_ptr()
        ctx.process_batch_device(d_in[0].data_ptr(), n, w, h, params, o_it, streams[0].cuda_stream)
        torch.cuda.synchronize()
        grown = float((d_out[0] > 0).float().mean().item())
        for i in range(8):
            ctx.process_batch_device(d_in[i % series].data_ptr(), n, w, h, params, outs[i % 4], streams[i % 4].cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(streams[0])
        for i in range(steps):
            ctx.process_batch_device(d_in[i % series].data_ptr(), n, w, h, params, outs[i % 4], streams[i % 4].cuda_stream)
        for s in streams[1:]:
            streams[0].wait_stream(s)
        e1.record(streams[0])
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    print(f"{name}: {n} slices {w}x{h} per step, {steps} steps, {series} series resident ({host.nbytes >> 20} MiB): "
          f"{ms / steps * 1e3:.1f} us/step, {n * steps / (ms * 1e-3):,.0f} slices/s, {w * h * n * steps / (ms * 1e-3) / 1e9:.1f} Gpx/s; "
          f"fill passes mean {float(d_it.float().mean()):.1f} max {int(d_it.max())}; dilated mask covers {100 * grown:.1f} % of the slice")


run("C2 default", 512, 512, 24, "default", 12)
run("C5 low contrast", 512, 512, 24, "lowcontrast", 12)
run("C4 1024x1024, median 5, dilation 7", 1024, 1024, 12, "default", 8, median_window=5, dilate_size=7)
