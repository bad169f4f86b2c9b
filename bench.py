#!/usr/bin/env python3
"""bench.py — slices/sec of the full per-slice pipeline (BASELINE.json metric) on N B200s.

A step = one pass of the hot path (K1 fused preprocess -> K23 flood fill + morphology) over one batch of synthetic
slices per GPU.  `--config` picks the BASELINE.json configuration (default c2, the one the metric is quoted on):

  c2  one synthetic patient series: 24 slices 512x512 int16, default parameters          (weak scaling)
  c3  cohort of 20 patients x 24 slices 512x512, contiguous patient ranges per rank; a step = the whole cohort
      (480 slices over all ranks), the slowest rank decides                              (strong scaling)
  c4  large matrix: 12 slices 1024x1024 per step, median window 5, dilation size 7       (weak)
  c5  worst-case region growing: 24 low-contrast 512x512 slices per step                 (weak)

Keys of the JSON line (one line, rank 0):
  value          device-resident throughput: inputs already in HBM, CUDA events on the launch stream, best of
                 REGIONS timed regions of exactly --steps steps each (all regions listed in `regions_ms`)
  e2e            same metric through the host-buffer entry point with pinned HOST buffers (H2D + kernels + D2H inside
                 the timed region); `e2e.link_*` = a raw pinned-copy probe of the same pattern run on all ranks at
                 once just before; `e2e.with_io` = the C++ driver from DICOM files to masks and to JPEGs on all GPUs
  roofline       dominant kernel (K1) against the measured HBM copy peak, plus the ALU-pipe ceiling that binds it
  parity_checked the outputs of one timed step and of one e2e step, compared bit for bit with the CPU oracle
                 (outside every clock); a mismatch makes the run fail
  cpu_baseline   the CPU oracle (a port of FAST's per-filter semantics; the reference's own FAST build is
                 unavailable) in the reference's OpenMP batch structure, timed on this box's cores

`--impl reference` times that CPU path alone, in the same JSON shape.
"""
import argparse
import ctypes as C
import hashlib
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "nm03-capstone-project_b200"

METRIC = "slices/sec (512x512 int16, full pipeline)"
UNIT = "slices/s"
REGIONS = 5                  # timed regions of --steps steps each; the best one is `value` (best of 5)
L2_BYTES = 126 << 20

CONFIGS = {
    "c2": dict(w=512, h=512, slices=24, kind="default", params={}, scaling="weak",
               workload="one synthetic patient series: 24 slices 512x512 int16, default median/sharpen/dilation params"),
    "c3": dict(w=512, h=512, slices=24, kind="default", params={}, scaling="strong", patients=20,
               workload="synthetic cohort: 20 patients x 24 slices 512x512 int16, sharded by patient across the ranks"),
    "c4": dict(w=1024, h=1024, slices=12, kind="default", params=dict(median_window=5, dilate_size=7), scaling="weak",
               workload="large-matrix stress: 1024x1024 int16 slices (12 per step per GPU), median window 5, dilation size 7 (radius 3)"),
    "c5": dict(w=512, h=512, slices=24, kind="lowcontrast", params={}, scaling="weak",
               workload="worst-case region growing: low-contrast phantom, grown region > 50 % of each 512x512 slice (24 per step per GPU)"),
}


def host_threads():
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1 to its workers, which
    is not what the reference's `omp_set_num_threads(16)` CPU run should be limited to)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def bind_to_gpu_numa_node(torch, index):
    """Restrict this rank (and the feeder threads it starts later) to the CPUs of the NUMA node its GPU hangs off, so the
    pinned staging buffers are first-touched in the memory next to that GPU's PCIe root.  Returns a note for `config`."""
    if os.environ.get("NM03_BENCH_NUMA", "1") == "0":
        return "off"
    try:
        p = torch.cuda.get_device_properties(index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return "single node"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return f"node {node} has no cpu in this process's mask"
        os.sched_setaffinity(0, allowed)
        return f"node {node}, {len(allowed)} cpus"
    except (OSError, AttributeError, ValueError) as e:
        return f"unavailable ({type(e).__name__})"


def load_oracle():
    """CPU oracle — used ONLY for the cpu_baseline / --impl reference legs and the parity check after the clocks."""
    so = os.path.join(ROOT, "oracle", "libnm03_oracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)
    lib = C.CDLL(so)
    lib.nm03o_run_batch_omp.restype = C.c_int
    lib.nm03o_max_threads.restype = C.c_int
    return lib


def oracle_params(cfg):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common as cm
    return cm.oparams(**cfg["params"])


def oracle_masks(oracle, series, cfg, threads):
    """Dilated masks of an (n, h, w) int16 host array from the oracle (OpenMP batch-of-25 structure)."""
    import numpy as np
    series = np.ascontiguousarray(series)
    n, h, w = series.shape
    out = np.empty((n, h, w), np.uint8)
    p = oracle_params(cfg)
    ok = oracle.nm03o_run_batch_omp(series.ctypes.data_as(C.c_void_p), n, w, h, C.byref(p), 25, threads,
                                    out.ctypes.data_as(C.c_void_p))
    assert ok == n, f"oracle processed {ok} of {n} slices"
    return out


def numpy_series(seed, cfg):
    """One synthetic T1+C-like series on the host (same phantom family as tests/common.py)."""
    import numpy as np
    w, h, n = cfg["w"], cfg["h"], cfg["slices"]
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    out = np.empty((n, h, w), np.int16)
    for s in range(n):
        e = ((xx - w / 2) / (0.42 * w)) ** 2 + ((yy - h / 2) / (0.48 * h)) ** 2
        img = np.maximum(rng.normal(40, 15, (h, w)), 0)
        head = e <= 1.0
        if cfg["kind"] == "lowcontrast":
            img[head] = rng.normal(1500, 150, (h, w))[head]
        else:
            img[head] = rng.normal(1000, 60, (h, w))[head]
            ring = (e > 0.88) & head
            img[ring] = rng.normal(600, 30, (h, w))[ring]
            for b in range(1 + s % 3):
                bx = w / 2 + (b - 1) * 0.14 * w + rng.uniform(-10, 10) * w / 512
                by = h / 2 + rng.uniform(-30, 30) * h / 512
                r = (0.06 + 0.02 * ((s % 8) / 8.0)) * w * (1 + 0.3 * b)
                blob = (xx - bx) ** 2 + (yy - by) ** 2 <= r * r
                img[blob] = rng.normal(1600, 120, (h, w))[blob]
        out[s] = np.clip(np.rint(img), 0, 4095).astype(np.int16)
    return out


def torch_series(torch, n_series, device, seed, cfg=None):
    """n_series synthetic series generated on the GPU (same structure as numpy_series)."""
    cfg = cfg or CONFIGS["c2"]
    W, H, SL = cfg["w"], cfg["h"], cfg["slices"]
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float32),
                            torch.arange(W, device=device, dtype=torch.float32), indexing="ij")
    e = ((xx - W / 2) / (0.42 * W)) ** 2 + ((yy - H / 2) / (0.48 * H)) ** 2
    head, ring = e <= 1.0, (e > 0.88) & (e <= 1.0)
    out = torch.empty((n_series, SL, H, W), dtype=torch.int16, device=device)
    for p in range(n_series):
        nrm = lambda m, s: torch.randn((SL, H, W), generator=g, device=device) * s + m  # noqa: E731
        img = torch.clamp(nrm(40, 15), min=0)
        if cfg["kind"] == "lowcontrast":
            img = torch.where(head, nrm(1500, 150), img)
        else:
            img = torch.where(head, nrm(1000, 60), img)
            img = torch.where(ring, nrm(600, 30), img)
            blobs = nrm(1600, 120)
            for s in range(SL):
                for b in range(1 + (p + s) % 3):
                    u = torch.rand(2, generator=g, device=device)
                    bx = W / 2 + (b - 1) * 0.14 * W + (u[0] * 20 - 10) * W / 512
                    by = H / 2 + (u[1] * 60 - 30) * H / 512
                    r = (0.06 + 0.02 * ((s % 8) / 8.0)) * W * (1 + 0.3 * b)
                    m = (xx - bx) ** 2 + (yy - by) ** 2 <= r * r
                    img[s] = torch.where(m, blobs[s], img[s])
        out[p] = torch.clamp(torch.round(img), 0, 4095).to(torch.int16)
    return out


def time_cpu(oracle, series, cfg, threads, min_seconds, max_reps):
    """Oracle in the reference's OpenMP batch structure (main_parallel.cpp:330-347) on one series."""
    import numpy as np
    p = oracle_params(cfg)
    n, h, w = series.shape
    out = np.empty(series.shape, np.uint8)
    reps, t0 = 0, time.perf_counter()
    while True:
        ok = oracle.nm03o_run_batch_omp(series.ctypes.data_as(C.c_void_p), n, w, h, C.byref(p), 25, threads,
                                        out.ctypes.data_as(C.c_void_p))
        assert ok == n
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= min_seconds or reps >= max_reps:
            return reps * n / dt, reps, dt


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed regions (device-resident, then e2e) run."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))  # noqa: E702
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def kernel_sources_sha():
    """Hash of the CUDA sources the kernels are built from: an ncu summary is only quoted when it was captured from these."""
    hsh = hashlib.sha256()
    for name in ("nm03_kernels.cu", "nm03_k1.cuh", "nm03_k1_w.cu", "nm03_device.h", "median_net.inc"):
        with open(os.path.join(ROOT, PKG, "csrc", name), "rb") as f:
            hsh.update(f.read())
    return hsh.hexdigest()[:16]


def ncu_capture_for_current_kernels():
    """(summary of K1, file name) from the newest profiles/*_ncu_summary.json stamped with the hash of the CURRENT kernel
    sources, else (None, reason).  A capture of other sources says nothing about the kernels that were just timed."""
    import glob
    sha = kernel_sources_sha()
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_summary.json")), reverse=True):
        try:
            summ = json.load(open(path))
        except (OSError, ValueError):
            continue
        if summ.get("kernel_sources_sha") != sha:
            continue
        k1 = next((v for k, v in summ.items() if "k1_preprocess" in k), None)
        if k1:
            return k1, "profiles/" + os.path.basename(path)
    return None, f"no ncu capture of the current kernel sources (sha {sha}) under profiles/"


def link_probe(torch, dev, h2d_bytes, d2h_bytes, seconds, barrier):
    """Raw pinned host<->device copies in the e2e pattern (one H2D and one D2H stream, whole-step transfers), all ranks
    at once: what the box's host side can deliver to this rank while the others copy too.  GB/s per direction."""
    h_src = torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory()
    h_dst = torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(h2d_bytes, dtype=torch.uint8, device=dev)
    d_b = torch.empty(d2h_bytes, dtype=torch.uint8, device=dev)
    s_in, s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    for _ in range(3):
        with torch.cuda.stream(s_in):
            d_a.copy_(h_src, non_blocking=True)
        with torch.cuda.stream(s_out):
            h_dst.copy_(d_b, non_blocking=True)
    # both copy engines stay busy the whole time: the next burst is queued before the previous one is waited for
    BURST = 16
    best = (0.0, 0.0)
    for _ in range(3):                       # three rounds of seconds/3, the best one counts (the first also warms the path)
        barrier()
        reps, t0 = 0, time.perf_counter()
        ev_prev = None
        while time.perf_counter() - t0 < seconds / 3:
            for _ in range(BURST):
                with torch.cuda.stream(s_in):
                    d_a.copy_(h_src, non_blocking=True)
                with torch.cuda.stream(s_out):
                    h_dst.copy_(d_b, non_blocking=True)
            ev = (s_in.record_event(), s_out.record_event())
            if ev_prev is not None:
                ev_prev[0].synchronize()
                ev_prev[1].synchronize()
                reps += BURST
            ev_prev = ev
        ev_prev[0].synchronize()
        ev_prev[1].synchronize()
        reps += BURST
        dt = time.perf_counter() - t0
        if reps * (h2d_bytes + d2h_bytes) / dt > sum(best) * 1e9:
            best = (reps * h2d_bytes / dt / 1e9, reps * d2h_bytes / dt / 1e9)
    return best


def driver_io_figure(n_gpus):
    """north_star's own end-to-end: the C++ batch driver from DICOM files (parse on host threads -> pinned staging ->
    GPU -> masks, and -> 2 JPEGs per slice) on every GPU of the job, on the C3 cohort and on a cohort long enough to
    amortise start-up; beside it the reference's OWN img_processing_parallel, compiled unchanged against the shim."""
    import tempfile
    host = os.path.join(ROOT, PKG, "host")
    drv = os.path.join(host, "nm03_driver")
    if not os.path.exists(drv):
        return None
    gpus = ",".join(str(i) for i in range(n_gpus))
    try:
        shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None   # memory-backed: no disk in the figure
        if shm:
            st = os.statvfs(shm)
            if st.f_bavail * st.f_frsize < (2 << 30):
                shm = None                       # a small tmpfs (container default 64 MiB): the ordinary temp dir instead
        free = os.statvfs(shm or tempfile.gettempdir())
        big = 120 if free.f_bavail * free.f_frsize > (6 << 30) else 40
        out = {"gpus": n_gpus, "host_cores": host_threads()}
        with tempfile.TemporaryDirectory(dir=shm) as td:
            for label, patients in (("c3_480_slices", 20), (f"long_{big * 24}_slices", big)):
                data = f"{td}/{label}/Brain-Tumor-Progression/T1-Post-Combined-P001-P020"
                subprocess.check_call([drv, "--make-dataset", data, "--patients", str(patients), "--slices", "24", "--size", "512"],
                                      stdout=subprocess.DEVNULL)
                rec = {}
                for name, extra in (("dicom_to_mask", ["--no-export"]), ("dicom_to_jpeg", [])):
                    best = None
                    for _ in range(2):   # the first process pays for cold files: best of two
                        r = subprocess.run([drv, "--data", data, "--out", f"{td}/o", "--gpus", gpus] + extra,
                                           capture_output=True, text=True, timeout=600)
                        one = json.loads(r.stdout.strip().splitlines()[-1])
                        if best is None or one["slices_per_sec"] > best["slices_per_sec"]:
                            best = one
                    rec[name + "_slices_per_sec"] = best["slices_per_sec"]
                    rec[name + "_startup_seconds"] = best.get("startup_seconds")
                out[label] = rec
                ref = os.path.join(host, "_refbuild", "img_processing_parallel")
                if label.startswith("c3") and os.path.exists(ref):
                    # the reference's main_parallel.cpp as it is (16 OpenMP threads, one slice per call, serial export,
                    # main_parallel.cpp:336-346,401) on the same files, through the FAST-shaped shim; whole process timed
                    work = f"{td}/refwork"
                    os.makedirs(work, exist_ok=True)
                    env = dict(os.environ, NM03_TEST_DATA_PATH=f"{td}/{label}/", OMP_NUM_THREADS="16")
                    t0 = time.perf_counter()
                    r = subprocess.run([ref], cwd=work, env=env, capture_output=True, text=True, timeout=900)
                    dt = time.perf_counter() - t0
                    done = r.stdout.count("Successfully processed 24/24 images")
                    out["reference_driver_on_shim"] = {"slices_per_sec": done * 24 / dt, "patients_ok": done, "seconds": dt,
                                                       "note": "reference main_parallel.cpp compiled unchanged against host/compat + libnm03.so, 1 GPU, wall time of the whole process incl. start-up and JPEG export"}
        out["sample"] = ("DICOM files in " + ("a memory-backed temp dir" if shm else "the temp dir") + ", timed after context start-up, best of two runs each")
        return out
    except Exception as e:  # informational only
        return {"error": f"{type(e).__name__}: {str(e)[:200]}"}


def run_reference(args, rank, cfg):
    """--impl reference: the reference's CPU implementation of the path.  FAST itself cannot be built
    here, so this is the oracle port in main_parallel.cpp's OpenMP structure."""
    if rank != 0:
        return
    oracle = load_oracle()
    threads = host_threads()
    series = numpy_series(1, cfg)
    n = series.shape[0]
    for _ in range(args.warmup):
        time_cpu(oracle, series, cfg, threads, 0, 1)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        time_cpu(oracle, series, cfg, threads, 0, 1)
    dt = time.perf_counter() - t0
    val = args.steps * n / dt
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000 * dt / args.steps, "higher_is_better": True, "scaling": cfg["scaling"],
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["workload"], "note": "CPU oracle port of FAST's per-filter semantics in the reference's "
                   "OpenMP batch-of-25 structure (main_parallel.cpp:330-347); FAST itself is unavailable on this box"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{args.steps} x one {n}-slice {cfg['w']}x{cfg['h']} series"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU time spent on the cpu_baseline sample")
    ap.add_argument("--no-io", action="store_true", help="skip the DICOM->JPEG driver figure (e2e.with_io)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    cfg = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank, cfg)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    full_mask = os.sched_getaffinity(0)
    numa_note = bind_to_gpu_numa_node(torch, local_rank) if world > 1 else "not needed at 1 GPU"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    nm03 = importlib.import_module(PKG)
    ctx = nm03.Context([local_rank])
    params = nm03.default_params(**cfg["params"])
    W, H, SLICES = cfg["w"], cfg["h"], cfg["slices"]
    series_bytes = SLICES * W * H * 2

    # ---- resident inputs (larger than L2), outputs ------------------------------------------
    # c2/c4/c5: one series per step, N_SERIES distinct ones rotated.  c3: this rank's contiguous patient range is one
    # step (one call per patient); enough distinct copies of the range are rotated to exceed L2.
    if args.config == "c3":
        lo, hi = nm03.shard_range(cfg["patients"], world, rank)
        calls_per_step = hi - lo
        variants = max(2, -(-int(2.4 * L2_BYTES) // max(1, calls_per_step * series_bytes)))
        N_SERIES = variants * calls_per_step
        units_per_step_job = cfg["patients"] * SLICES            # the whole cohort, all ranks together
    else:
        calls_per_step = 1
        N_SERIES = max(4, -(-int(2.4 * L2_BYTES) // series_bytes))
        units_per_step_job = world * SLICES
    d_in = torch_series(torch, N_SERIES, dev, 1234 + rank, cfg)
    # Device-resident steps are issued round-robin on a few streams (a resident caller with more than one series in
    # flight): the library keeps a scratch slot per calling stream, so the latency-bound fill/morphology kernel of
    # one series runs under the ALU-bound K1 of the next.  NM03_BENCH_STREAMS=1 gives the single-stream figure.
    n_streams = max(1, min(4, int(os.environ.get("NM03_BENCH_STREAMS", "4"))))
    args.warmup = max(args.warmup, n_streams)   # every stream (and its scratch slot in the library) is warm before the clock starts
    streams = [torch.cuda.Stream(device=dev) for _ in range(n_streams)]
    stream = streams[0]
    d_out = torch.empty((max(2, n_streams), SLICES, H, W), dtype=torch.uint8, device=dev)   # one per stream in flight
    outs = []
    for i in range(d_out.shape[0]):
        o = nm03.Outputs()
        o.dilated = d_out[i].data_ptr()
        outs.append(o)

    # everything a call needs is looked up once: the issue loop below must stay well under the 45 us a step takes on the
    # GPU (at 8 ranks on one 32-vCPU host it is the loop, not the GPU, that a careless harness would time)
    in_ptrs = [C.c_void_p(d_in[j].data_ptr()) for j in range(N_SERIES)]
    stream_handles = [C.c_void_p(s.cuda_stream) for s in streams]
    out_refs = [C.byref(o) for o in outs]
    params_ref = C.byref(params)
    raw_call, ctx_handle = nm03.lib().nm03_process_batch_device, ctx._h

    def call(j, n_str):   # one library call: one series
        rc = raw_call(ctx_handle, 0, in_ptrs[j % N_SERIES], SLICES, W, H, params_ref, None, 0, out_refs[j % len(outs)],
                      stream_handles[j % n_str])
        if rc:
            raise nm03.Nm03Error(rc, nm03.last_error())

    def step(i, n_str):
        for c in range(calls_per_step):
            call(i * calls_per_step + c, n_str)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    issue_s = [0.0]
    next_step = [0]

    def timed_region(steps, n_str):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)                      # every stream is idle here (barrier synchronised the device)
        t_issue = time.perf_counter()
        for _ in range(steps):
            step(next_step[0], n_str)
            next_step[0] += 1
        issue_s[0] = time.perf_counter() - t_issue   # host time spent queueing the steps (must stay below the device time)
        for other in streams[1:n_str]:
            stream.wait_stream(other)          # e1 fires after the last step of every stream
        e1.record(stream)
        barrier()
        return e0.elapsed_time(e1)

    # the clock sampler runs through every timed region; it is started before the warm-up and given time to deliver its
    # first line, because nvidia-smi's start-up (NVML initialisation, device enumeration) holds driver locks that stall
    # kernel launches of every rank for a few hundred milliseconds
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_wait = time.perf_counter()
    while sampler.proc and not sampler.lines and time.perf_counter() - t_wait < 5.0:
        time.sleep(0.05)
    for _ in range(args.warmup):
        step(next_step[0], n_streams)
        next_step[0] += 1
    barrier()
    assert int(d_out[(next_step[0] * calls_per_step - 1) % len(outs)].sum().item()) > 0, "pipeline produced empty masks"
    launches0 = ctx.kernel_launches()
    regions = [timed_region(args.steps, n_streams) for _ in range(REGIONS)]
    launches = (ctx.kernel_launches() - launches0) // REGIONS   # per region of --steps steps
    issue_ms_per_step = issue_s[0] * 1e3 / args.steps

    # iterations-to-convergence of the region growing on this workload (outside every timed region)
    d_it = torch.zeros(SLICES, dtype=torch.int32, device=dev)
    o_it = nm03.Outputs()
    o_it.dilated = d_out[0].data_ptr()
    o_it.rg_iterations = d_it.data_ptr()
    ctx.process_batch_device(d_in[0].data_ptr(), SLICES, W, H, params, o_it, stream.cuda_stream)
    torch.cuda.synchronize()
    rg_passes = {"mean": float(d_it.float().mean().item()), "max": int(d_it.max().item()),
                 "grown_fraction_of_slice": float((d_out[0] > 0).float().mean().item()),
                 "note": "fill passes (down+up sweeps over the slice) until no pixel changes, per slice of one series"}

    # Per-kernel durations for the roofline: the same steps once more on ONE stream, where a kernel's CUDA-event
    # duration is its own (with two streams the events of one step also span the neighbouring step's kernels).
    # Events are recorded around each kernel on every 8th call only: an event between K1 and K23 defeats their
    # programmatic overlap (~6 % per step).
    PROFILE_EVERY = 8
    serial_steps = min(args.steps, 1000)
    real_call = call
    n_calls = [0]

    def call(j, n_str):   # noqa: F811 - the profiled variant of the same call
        ctx.profile_enable(n_calls[0] % PROFILE_EVERY == 0)   # the first call of the region is always sampled
        n_calls[0] += 1
        real_call(j, n_str)

    ctx.profile_work()                                   # reset the sorted-window counter
    launches_s0 = ctx.kernel_launches()
    serial_ms = timed_region(serial_steps, 1)
    k1_ms, k23_ms, calls = ctx.profile_read()
    sorted_runs = ctx.profile_work()                     # median work items K1 actually sorted in that region
    k1_launches_serial = (ctx.kernel_launches() - launches_s0) // 2
    # and once more as in the headline region (all streams), for the record: there the events around a kernel also span
    # whatever the neighbouring steps run at the same time
    n_calls[0] = 0
    timed_region(serial_steps, n_streams)
    k1_ov_ms, k23_ov_ms, calls_ov = ctx.profile_read()
    ctx.profile_enable(False)
    call = real_call

    # ---- parity of what was just timed (outside every clock) ------------------------------------
    # one series through the same call, on the stream and output buffer the timed steps used; the oracle gets the same bytes
    j_chk = (next_step[0] * calls_per_step - 1) % N_SERIES
    torch.cuda.synchronize()
    ctx.process_batch_device(d_in[j_chk].data_ptr(), SLICES, W, H, params, outs[0], stream.cuda_stream)
    torch.cuda.synchronize()
    chk_in = d_in[j_chk].cpu().numpy()
    chk_dev = d_out[0].cpu().numpy()

    # ---- e2e: pinned host buffers through the public entry point ------------------------------
    # Every step copies the series in and the masks out inside the timed region.  One caller thread keeps several
    # batches in flight through the ticketed asynchronous entry point (nm03_process_batch_async / nm03_wait);
    # NM03_BENCH_FEEDERS=k instead runs k threads over the blocking nm03_process_batch (the reference's threading contract).
    n_feed = int(os.environ.get("NM03_BENCH_FEEDERS", "0"))
    depth = 3
    n_hin = 4
    h_in = ctx.host_alloc((n_hin, SLICES, H, W), np.int16)
    for i in range(n_hin):
        h_in[i] = d_in[i].cpu().numpy()
    n_hout = max(n_feed, depth)
    h_outs = [ctx.host_alloc((SLICES, H, W), np.uint8) for _ in range(n_hout)]
    e2e_steps = max(60, min(args.steps, 400)) * calls_per_step   # at least 60 batches: with three in flight a shorter run is mostly ramp and drain
    e2e_params = nm03.default_params(**cfg["params"])

    def feeder(tid, count):
        res = {"dilated": h_outs[tid]}
        for i in range(count):
            ctx.process_batch(h_in[(tid + n_feed * i) % n_hin], e2e_params, want=("dilated",), out=res)

    def run_e2e(total):
        """returns the index of the input the LAST completed batch of output buffer 0 was computed from"""
        if n_feed > 0:
            ths = [threading.Thread(target=feeder, args=(t, total // n_feed)) for t in range(n_feed)]
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            return (0 + n_feed * (total // n_feed - 1)) % n_hin
        tickets = [None] * depth
        src0 = 0
        for i in range(total):
            b = i % depth
            if tickets[b] is not None:
                ctx.wait(tickets[b])
            tickets[b] = ctx.process_batch_async(h_in[i % n_hin], e2e_params, out_bufs[b])
            if b == 0:
                src0 = i % n_hin
        for t in tickets:
            if t is not None:
                ctx.wait(t)
        return src0

    out_bufs = [{"dilated": h} for h in h_outs]

    run_e2e(8 * max(n_feed, depth))   # warm-up: every stream slot of the library has sized its staging buffers before the clock starts
    h2d_gbs, d2h_gbs = link_probe(torch, dev, series_bytes, SLICES * W * H, 1.5, barrier)
    barrier()
    t0 = time.perf_counter()
    e2e_src0 = run_e2e(e2e_steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if n_feed > 0:
        e2e_steps = (e2e_steps // n_feed) * n_feed
    # the same once more with the result fetched bit-packed (nm03_outputs.dilated_bits): 1/8 of the D2H bytes
    bits_s = None
    if n_feed == 0:
        h_bits = [ctx.host_alloc((SLICES, H, (W + 31) // 32), np.uint32) for _ in range(depth)]
        out_bufs = [{"dilated_bits": h} for h in h_bits]
        run_e2e(2 * depth)
        barrier()
        t0 = time.perf_counter()
        bits_src0 = run_e2e(e2e_steps)
        torch.cuda.synchronize()
        bits_s = time.perf_counter() - t0
        bits_ok = bool((nm03.unpack_bits(h_bits[0], W) == np.array(h_outs[0])).all()) if bits_src0 == e2e_src0 else None
    clocks = sampler.stop()
    assert all(int(h.sum()) > 0 for h in h_outs)

    # ---- max over ranks ------------------------------------------------------------------------
    t = torch.tensor(regions + [e2e_s * 1000.0, k1_ms, k23_ms, serial_ms, (bits_s or 0.0) * 1000.0], dtype=torch.float64, device=dev)
    link = torch.tensor([h2d_gbs, d2h_gbs], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(link, op=dist.ReduceOp.SUM)
        tl = torch.tensor([float(launches)], dtype=torch.float64, device=dev)
        dist.all_reduce(tl, op=dist.ReduceOp.SUM)
        launches = int(tl.item())
    tt = [float(x) for x in t.tolist()]
    regions, (e2e_ms, k1_ms, k23_ms, serial_ms, bits_ms) = tt[:REGIONS], tt[REGIONS:]
    ms = min(regions)

    # ---- parity check (every rank checks its own outputs; a mismatch anywhere fails the job) ------
    oracle = load_oracle()
    os.sched_setaffinity(0, full_mask)
    threads = host_threads()
    exp_dev = oracle_masks(oracle, chk_in, cfg, threads)
    exp_e2e = oracle_masks(oracle, np.array(h_in[e2e_src0]), cfg, threads)
    mism = int((exp_dev != chk_dev).reshape(SLICES, -1).any(axis=1).sum()) + \
        int((exp_e2e != h_outs[0]).reshape(SLICES, -1).any(axis=1).sum())
    mt = torch.tensor([float(mism)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(mt, op=dist.ReduceOp.SUM)
    mism_all = int(mt.item())

    with_io = None
    want_io = not args.no_io and os.environ.get("NM03_BENCH_IO", "1") != "0"
    if want_io and world > 1:
        # the other ranks SLEEP while rank 0 runs the driver on every GPU (a NCCL barrier would keep one core per rank
        # spinning next to the driver's host threads)
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            with_io = driver_io_figure(world)
            store.set("nm03_io_done", "1")
        else:
            while not store.check(["nm03_io_done"]):
                time.sleep(0.2)
    elif want_io:
        with_io = driver_io_figure(world)
    barrier()

    if rank == 0:
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak, peak_src = float(peaks["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        except (OSError, KeyError, ValueError):
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        # K1 in this configuration must read 2 B/px of int16 and write the 1-bit/px predicate plane
        k1_bytes = SLICES * W * H * (2.0 + 1.0 / 8.0)
        # DRAM traffic of one K1 launch: only from an `ncu --set full` capture of exactly these kernel sources
        k1s, traffic_src = ncu_capture_for_current_kernels()
        traffic = (k1s["dram_read_bytes"] + k1s["dram_write_bytes"]) if (k1s and args.config == "c2") else None
        if k1s and args.config != "c2":
            traffic_src = "captures are made on c2 only"
        k1_avg_ms = k1_ms / max(calls, 1)
        achieved = k1_bytes / (k1_avg_ms * 1e-3) / 1e9
        step_ms = ms / args.steps
        # the unit that binds K1: packed min/max instructions of the median networks on the 64-lane/clk/SM ALU pipe
        # (tools/ubench_minmax.cu).  sorted_runs is counted by the kernel itself (windows that take the clip-floor
        # shortcut are not sorted); instructions per run come from the generated network.
        ops_per_run = nm03.median_network_ops(params.median_window)
        sm_clock_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        alu_peak = 64.0 * 148 * sm_clock_hz                       # lane-instructions per second
        minmax_per_launch = sorted_runs * ops_per_run / max(k1_launches_serial, 1)
        alu = {"bound": "alu-pipe (packed s16x2 min/max)", "minmax_lane_ops_per_launch": minmax_per_launch,
               "achieved": minmax_per_launch / (k1_avg_ms * 1e-3) / 1e12, "peak": alu_peak / 1e12, "unit": "T lane-ops/s",
               "frac": minmax_per_launch / (k1_avg_ms * 1e-3) / alu_peak,
               "sorted_runs_per_launch": sorted_runs / max(k1_launches_serial, 1), "ops_per_run": ops_per_run,
               "peak_source": "64 lanes/clk/SM x 148 SMs x the SM clock sampled during this run",
               "note": "min/max instructions of the median networks only; the other ALU-pipe instructions of K1 "
                       "(permutes, compares, address arithmetic) share the same pipe and are not counted as useful work"}
        if k1s:
            alu["ncu"] = {"alu_pipe_pct_of_peak_active": float(k1s["alu_pipe_pct_active"]),
                          "issue_slots_pct_active": float(k1s["issue_active_pct"]),
                          "warp_instructions_per_launch": float(k1s["warp_inst"]), "source": traffic_src}
        cpu_val, reps, cpu_dt = time_cpu(oracle, np.array(h_in[0]), cfg, threads, args.cpu_seconds, 1000)
        e2e_val = world * SLICES * e2e_steps / (e2e_ms * 1e-3)     # every rank pushes e2e_steps batches of SLICES slices
        e2e_gbs = e2e_val * W * H * 3 / 1e9
        link_h2d, link_d2h = [float(x) for x in link.tolist()]
        line = {
            "metric": METRIC, "value": units_per_step_job * args.steps / (ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms,
            "regions_ms": regions, "regions_spread": (max(regions) - min(regions)) / min(regions),
            "host_issue_ms_per_step": issue_ms_per_step, "higher_is_better": True,
            "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "s16x2 min/max + f32", "data": "synthetic",
            "config": {"workload": cfg["workload"], "name": args.config, "slices_per_call": SLICES,
                       "calls_per_step_per_gpu": calls_per_step, "parallelism": f"patient-range x{world}",
                       "l2": f"rotating {N_SERIES} distinct resident series ({N_SERIES * series_bytes >> 20} MiB > 126 MB L2)",
                       "timing": f"best of {REGIONS} regions of {args.steps} steps each, barrier + synchronize on both sides, max over ranks per region",
                       "outputs": "dilated uint8 mask", "resident_streams": n_streams,
                       "jpeg_export": "host stage; in e2e.with_io", "rank_numa_binding": numa_note},
            "e2e": {"value": e2e_val, "unit": UNIT,
                    "h2d_bytes_per_step": SLICES * W * H * 2 * calls_per_step, "d2h_bytes_per_step": SLICES * W * H * calls_per_step,
                    "steps": e2e_steps // calls_per_step,
                    "api": (f"nm03_process_batch (pinned host in/out), {n_feed} feeder threads per GPU" if n_feed > 0 else
                            f"nm03_process_batch_async + nm03_wait (pinned host in/out), ONE caller thread per GPU, {depth} batches in flight"),
                    "link_h2d_gbs": link_h2d, "link_d2h_gbs": link_d2h, "link_ceiling_gbs": link_h2d + link_d2h,
                    "moved_gbs": e2e_gbs, "frac_of_link": e2e_gbs / (link_h2d + link_d2h) if link_h2d + link_d2h > 0 else None,
                    "link_note": "raw pinned copies in the same pattern (whole-step H2D next to whole-step D2H) on all ranks at once, best of 3 rounds, summed over ranks; "
                                 "a ceiling for equal work per rank only roughly: on the 8-GPU box four GPUs get 16 and four 29 GB/s, and a rank that finishes early leaves its share to the others",
                    "bitpacked": None if not bits_ms else {
                        "value": world * SLICES * e2e_steps / (bits_ms * 1e-3), "unit": UNIT,
                        "d2h_bytes_per_step": SLICES * H * ((W + 31) // 32) * 4 * calls_per_step, "matches_byte_mask": bits_ok,
                        "note": "same call with nm03_outputs.dilated_bits instead of the uint8 mask (the contract's default stays uint8)"},
                    "with_io": with_io},
            "gpu_launches": launches,
            "clocks": clocks,
            "parity_checked": {"slices": 2 * SLICES * world, "mismatches": mism_all,
                               "against": "CPU oracle (nm03o_run_batch_omp) on the same input bytes: one device-resident call of the timed region and the last e2e batch, every rank"},
            "roofline": {"bound": "hbm", "kernel": f"k1_preprocess<{params.median_window},4>", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "kernel_sources_sha": kernel_sources_sha(),
                         "peak_source": peak_src,
                         "avg_launch_ms": k1_avg_ms, "bytes_per_launch": k1_bytes,
                         "frac_headline": k1_bytes * calls_per_step / (step_ms * 1e-3) / 1e9 / peak,
                         "frac_headline_note": "K1's algorithmic bytes over the WHOLE step time of the headline region (all streams, K23 included): a lower bound on K1's fraction in that regime",
                         "measured_on": f"{serial_steps} steps of the same workload on one stream, events on every "
                                        f"{PROFILE_EVERY}th call (the headline region runs {n_streams} streams, where "
                                        "kernels of neighbouring steps overlap)",
                         "single_stream_ms_per_step": serial_ms / serial_steps,
                         "avg_launch_ms_with_all_streams": k1_ov_ms / max(calls_ov, 1),
                         "k23_avg_launch_ms_with_all_streams": k23_ov_ms / max(calls_ov, 1),
                         "share_of_step": (k1_avg_ms * calls_per_step) / (serial_ms / serial_steps) if serial_ms > 0 else None,
                         "sampled_launches": calls,
                         "k23_avg_launch_ms": k23_ms / max(calls, 1),
                         "region_growing_passes": rg_passes,
                         "alu": alu,
                         "note": "K1 is bound by the integer min/max pipe (median networks), not HBM; see DESIGN.md"},
            "cpu_baseline": {"value": cpu_val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{reps} x one {SLICES}-slice {W}x{H} series ({cpu_dt:.1f} s), OpenMP batch-of-25"},
        }
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    if mism_all:
        raise SystemExit(f"PARITY FAILURE: {mism_all} slices differ from the CPU oracle")


if __name__ == "__main__":
    main()
