#!/usr/bin/env python3
"""Quick per-kernel timing of the device-resident pipeline (CUDA events inside the library).
NM03_LIB selects the build.  usage: k_time.py [n_slices] [size] [steps]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
W = H = int(sys.argv[2]) if len(sys.argv) > 2 else 512
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
nm03 = importlib.import_module("nm03-capstone-project_b200")
dev = torch.device("cuda", 0)
ctx = nm03.Context([0])
d_in = bench.torch_series(torch, 24, dev, 1234, dict(bench.CONFIGS['c2'], w=W, h=H))          # 24 series x 24 slices
flat = d_in.reshape(-1, H, W)
n_batches = flat.shape[0] // n
d_out = torch.empty((n, H, W), dtype=torch.uint8, device=dev)
o = nm03.Outputs(); o.dilated = d_out.data_ptr()
it = torch.zeros(n, dtype=torch.int32, device=dev); o.rg_iterations = it.data_ptr()
params = nm03.default_params(median_window=int(os.environ.get('K_TIME_WINDOW', '7')))
st = torch.cuda.Stream(device=dev)
for i in range(5):
    ctx.process_batch_device(flat[(i % n_batches) * n:].data_ptr(), n, W, H, params, o, st.cuda_stream)
torch.cuda.synchronize()
prof = os.environ.get('K_TIME_NOPROF') != '1'
ctx.profile_enable(prof)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for i in range(steps):
    ctx.process_batch_device(flat[(i % n_batches) * n:].data_ptr(), n, W, H, params, o, st.cuda_stream)
e1.record(st)
torch.cuda.synchronize()
k1, k23, calls = ctx.profile_read() if prof else (0.0, 0.0, 1)
calls = max(calls, 1)
print(f"{os.path.basename(os.environ.get('NM03_LIB', 'libnm03.so')):28s} n={n:4d} step={e0.elapsed_time(e1)/steps*1e3:8.1f} us  "
      f"k1={k1/calls*1e3:8.1f} us  k23={k23/calls*1e3:7.1f} us  -> {n*steps/(e0.elapsed_time(e1)*1e-3):10.0f} slices/s  "
      f"fill passes max={int(it.max())} mean={float(it.float().mean()):.1f}")
