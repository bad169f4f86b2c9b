"""nm03-capstone-project_b200 — B200-native per-slice brain-MRI pipeline (ctypes view of libnm03.so).

Host-side mirror of the C ABI in include/nm03.h for Python callers (tests, bench.py).  The product is
the CUDA library; this module only marshals numpy arrays / device pointers into it.  There is no
fallback of any kind: if libnm03.so is missing or no CUDA device is present, calls raise.

Reference path: processSingleImage(), src/sequential/main_sequential.cpp:170-272.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NM03_LIB", os.path.join(_HERE, "libnm03.so"))  # NM03_LIB: tuning builds

# every symbol include/nm03.h declares
EXPORTS = [
    "nm03_init", "nm03_shutdown", "nm03_last_error", "nm03_version", "nm03_gpu_count", "nm03_default_params",
    "nm03_make_seeds", "nm03_process_batch", "nm03_process_batch_device", "nm03_host_alloc", "nm03_host_free",
    "nm03_preprocess", "nm03_region_grow", "nm03_morphology", "nm03_kernel_launches",
    "nm03_profile_enable", "nm03_profile_read", "nm03_process_batch_device_noalloc", "nm03_device_count",
    "nm03_process_batch_async", "nm03_wait", "nm03_profile_work", "nm03_median_network_ops",
]

NOALLOC_MAX_PRED_WORDS = 256 * 1024  # NM03_NOALLOC_MAX_PRED_WORDS

OK = 0
ERR_INVALID, ERR_TOO_SMALL, ERR_PARAM, ERR_NO_SEEDS, ERR_SEED_OOB = -1, -2, -3, -4, -5
ERR_UNSUPPORTED, ERR_TOO_LARGE, ERR_CUDA, ERR_NO_DEVICE, ERR_NOMEM, ERR_BUSY = -6, -7, -8, -9, -10, -11


class Params(C.Structure):
    """nm03_params (include/nm03.h)."""
    _fields_ = [
        ("norm_low", C.c_float), ("norm_high", C.c_float), ("norm_min", C.c_float), ("norm_max", C.c_float),
        ("clip_min", C.c_float), ("clip_max", C.c_float),
        ("median_window", C.c_int),
        ("sharpen_gain", C.c_float), ("sharpen_sigma", C.c_float), ("sharpen_mask", C.c_int),
        ("rg_min", C.c_float), ("rg_max", C.c_float),
        ("dilate_size", C.c_int), ("erode_size", C.c_int),
        ("border_mode", C.c_int), ("sharpen_fma", C.c_int), ("rg_mode", C.c_int),
        ("rg_connectivity", C.c_int), ("rg_border_expand", C.c_int), ("se_shape", C.c_int),
        ("input_unsigned", C.c_int),
    ]


class Outputs(C.Structure):
    """nm03_outputs (include/nm03.h)."""
    _fields_ = [
        ("dilated", C.c_void_p), ("eroded", C.c_void_p), ("rg_mask", C.c_void_p),
        ("sharpened", C.c_void_p), ("median", C.c_void_p), ("rg_iterations", C.c_void_p),
        ("dilated_bits", C.c_void_p),
        ("export_blocks", C.c_void_p), ("export_nz", C.c_void_p), ("export_rq", C.c_void_p),
        ("export_seg_blocks", C.c_void_p), ("export_seg_nz", C.c_void_p), ("export_seg_style", C.c_void_p),
    ]


class ExportStyle(C.Structure):
    """nm03_export_style (include/nm03.h)."""
    _fields_ = [("background", C.c_uint8), ("interior", C.c_uint8), ("border", C.c_uint8), ("border_radius", C.c_int)]


class Nm03Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"nm03 error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """dlopen libnm03.so (built by `make -C nm03-capstone-project_b200` / __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it first (python -c 'import __graft_entry__ as g; g.build()');"
                               " there is no CPU or PyTorch fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.nm03_last_error.restype = C.c_char_p
        _lib.nm03_kernel_launches.restype = C.c_longlong
        _lib.nm03_kernel_launches.argtypes = [C.c_void_p, C.c_int]
        _lib.nm03_profile_enable.argtypes = [C.c_void_p, C.c_int, C.c_int]
        _lib.nm03_profile_read.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.nm03_shutdown.argtypes = [C.c_void_p]
        _lib.nm03_shutdown.restype = None
        _lib.nm03_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        _lib.nm03_host_free.argtypes = [C.c_void_p]
        _lib.nm03_process_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                            C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _lib.nm03_process_batch_async.argtypes = _lib.nm03_process_batch.argtypes
        _lib.nm03_wait.argtypes = [C.c_void_p]
        _lib.nm03_profile_work.restype = C.c_longlong
        _lib.nm03_profile_work.argtypes = [C.c_void_p, C.c_int]
        _lib.nm03_median_network_ops.argtypes = [C.c_int]
        _lib.nm03_process_batch_device.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                                   C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _lib.nm03_process_batch_device_noalloc.argtypes = _lib.nm03_process_batch_device.argtypes
        _lib.nm03_preprocess.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p]
        _lib.nm03_region_grow.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_int, C.c_void_p, C.c_void_p]
        _lib.nm03_morphology.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.c_void_p]
    return _lib


def last_error():
    return lib().nm03_last_error().decode()


def _check(rc):
    if rc != 0:
        raise Nm03Error(rc, last_error())


def default_params(**kw):
    p = Params()
    lib().nm03_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def median_network_ops(window):
    """min/max instructions one median work item of the preprocess kernel executes (nm03_median_network_ops)."""
    return lib().nm03_median_network_ops(int(window))


def _copy_params(params, dtype):
    """A private copy of the caller's nm03_params with input_unsigned set from the array's dtype (the caller's object is
    not touched, so it can be reused for int16 and uint16 input alike)."""
    q = Params()
    C.memmove(C.byref(q), C.byref(params if params is not None else default_params()), C.sizeof(Params))
    q.input_unsigned = 1 if dtype == np.uint16 else 0
    return q


class Ticket:
    """Handle of a batch queued with Context.process_batch_async; keeps the buffers alive until wait()."""

    def __init__(self, handle, keep, res):
        self._h, self._keep, self.res = handle, keep, res


def make_seeds(w, h):
    """Adaptive seeds of main_sequential.cpp:214-241 as an (n, 2) int32 array of (x, y)."""
    xy = np.zeros((512, 2), np.int32)
    n = lib().nm03_make_seeds(w, h, xy.ctypes.data_as(C.c_void_p), 512)
    if n < 0:
        raise Nm03Error(n, last_error())
    return xy[:n].copy()


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    """nm03_ctx: one worker per listed GPU."""

    def __init__(self, gpu_ids=(0,)):
        ids = (C.c_int * len(gpu_ids))(*gpu_ids)
        self._h = C.c_void_p()
        _check(lib().nm03_init(ids, len(gpu_ids), C.byref(self._h)))

    def close(self):
        if self._h:
            lib().nm03_shutdown(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def kernel_launches(self, gpu=0):
        return lib().nm03_kernel_launches(self._h, gpu)

    def profile_enable(self, on=True, gpu=0):
        _check(lib().nm03_profile_enable(self._h, gpu, int(on)))

    def profile_read(self, gpu=0):
        """(preprocess_ms, segment_ms, calls) summed since the last read."""
        a, b, n = C.c_double(0), C.c_double(0), C.c_longlong(0)
        _check(lib().nm03_profile_read(self._h, gpu, C.byref(a), C.byref(b), C.byref(n)))
        return a.value, b.value, n.value

    def profile_work(self, gpu=0):
        """nm03_profile_work: median work items the preprocess kernel sorted since the last call (waits for the GPU)."""
        v = lib().nm03_profile_work(self._h, gpu)
        if v < 0:
            raise Nm03Error(int(v), "nm03_profile_work failed")
        return int(v)

    def host_alloc(self, shape, dtype):
        """numpy array backed by pinned host memory (nm03_host_alloc)."""
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        _check(lib().nm03_host_alloc(C.byref(p), nbytes))
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dtype).reshape(shape)
        self._pinned = getattr(self, "_pinned", [])
        self._pinned.append(p)
        return arr

    # ---- whole pipeline, host buffers -------------------------------------------------------
    @staticmethod
    def _host_outputs(n, h, w, want, res):
        o = Outputs()
        for name in want:
            if name in ("export", "export_rq"):   # res["export_rq"]: 64 float32 from the caller (jpeg::make_tables / nm03h_jpeg_rq)
                if o.export_rq:
                    continue
                nb = (h // 8) * (w // 8)
                res["export_rq"] = np.ascontiguousarray(res["export_rq"], np.float32)
                res.setdefault("export_blocks", np.empty((n, nb, 64), np.int16))
                res.setdefault("export_nz", np.empty((n, nb), np.uint64))
                for k in ("export_blocks", "export_nz", "export_rq"):
                    setattr(o, k, res[k].ctypes.data)
                if "export_seg_style" in res:   # an ExportStyle: also the rendering of the result (needs "dilated_bits" in want)
                    res.setdefault("export_seg_blocks", np.empty((n, nb, 64), np.int16))
                    res.setdefault("export_seg_nz", np.empty((n, nb), np.uint64))
                    o.export_seg_blocks = res["export_seg_blocks"].ctypes.data
                    o.export_seg_nz = res["export_seg_nz"].ctypes.data
                    o.export_seg_style = C.addressof(res["export_seg_style"])
                continue
            if name in ("export_blocks", "export_nz", "export_seg_blocks", "export_seg_nz", "export_seg_style"):
                continue
            if name not in res:
                if name in ("sharpened", "median"):
                    res[name] = np.empty((n, h, w), np.float32)
                elif name == "rg_iterations":
                    res[name] = np.zeros(n, np.int32)
                elif name == "dilated_bits":
                    res[name] = np.empty((n, h, (w + 31) // 32), np.uint32)
                else:
                    res[name] = np.empty((n, h, w), np.uint8)
            setattr(o, name, res[name].ctypes.data)
        return o

    def process_batch(self, slices, params=None, seeds=None, want=("dilated",), gpu=0, out=None):
        """slices: (n, h, w) int16/uint16 array.  Returns dict of requested outputs."""
        slices = np.ascontiguousarray(slices)
        n, h, w = slices.shape
        params = _copy_params(params, slices.dtype)
        res = out if out is not None else {}
        o = self._host_outputs(n, h, w, want, res)
        status = np.zeros(n, np.int32)
        sp = None if seeds is None else np.ascontiguousarray(seeds, np.int32)
        rc = lib().nm03_process_batch(self._h, gpu, _p(slices), n, w, h, C.byref(params), _p(sp),
                                      0 if sp is None else len(sp), C.byref(o), _p(status))
        res["status"] = status
        _check(rc)
        return res

    def process_batch_async(self, slices, params=None, out=None, want=("dilated",), seeds=None, gpu=0):
        """nm03_process_batch_async: queue the batch and return a Ticket; `out` maps output names to (preferably pinned)
        arrays that receive the results, which are valid after wait(ticket).  Raises Nm03Error(ERR_BUSY) when all
        stream slots of the GPU hold unfinished tickets."""
        slices = np.ascontiguousarray(slices)
        n, h, w = slices.shape
        params = _copy_params(params, slices.dtype)
        res = out if out is not None else {}
        o = self._host_outputs(n, h, w, tuple(res.keys()) if out is not None else want, res)
        sp = None if seeds is None else np.ascontiguousarray(seeds, np.int32)
        t = C.c_void_p()
        _check(lib().nm03_process_batch_async(self._h, gpu, _p(slices), n, w, h, C.byref(params), _p(sp),
                                              0 if sp is None else len(sp), C.byref(o), C.byref(t)))
        return Ticket(t, (slices, sp, res), res)

    def wait(self, ticket):
        """nm03_wait: block until the ticket's outputs are in its buffers; returns them."""
        h, ticket._h = ticket._h, None
        if h is None:
            raise ValueError("ticket was already waited for")
        _check(lib().nm03_wait(h))
        return ticket.res

    # ---- whole pipeline, device pointers ----------------------------------------------------
    def process_batch_device(self, d_slices_ptr, n, w, h, params, d_out: Outputs, stream=0, seeds=None, gpu=0):
        sp = None if seeds is None else np.ascontiguousarray(seeds, np.int32)
        _check(lib().nm03_process_batch_device(self._h, gpu, C.c_void_p(d_slices_ptr), n, w, h, C.byref(params),
                                               _p(sp), 0 if sp is None else len(sp), C.byref(d_out),
                                               C.c_void_p(stream)))

    def process_batch_device_noalloc(self, d_slices_ptr, n, w, h, params, d_out: Outputs, stream=0, seeds=None, gpu=0):
        """nm03_process_batch_device_noalloc: same as process_batch_device, but the scratch bit-plane is
        the library's built-in 1 MiB array, so the call allocates nothing (see include/nm03.h)."""
        sp = None if seeds is None else np.ascontiguousarray(seeds, np.int32)
        _check(lib().nm03_process_batch_device_noalloc(self._h, gpu, C.c_void_p(d_slices_ptr), n, w, h,
                                                       C.byref(params), _p(sp), 0 if sp is None else len(sp),
                                                       C.byref(d_out), C.c_void_p(stream)))

    # ---- per-operator entry points ----------------------------------------------------------
    def preprocess(self, raw, params=None, want=("median", "sharpened", "predicate"), gpu=0):
        raw = np.ascontiguousarray(raw)
        h, w = raw.shape
        params = _copy_params(params, raw.dtype)
        med = np.empty((h, w), np.float32) if "median" in want else None
        sh = np.empty((h, w), np.float32) if "sharpened" in want else None
        pred = np.empty((h, (w + 31) // 32), np.uint32) if "predicate" in want else None
        _check(lib().nm03_preprocess(self._h, gpu, _p(raw), w, h, C.byref(params), _p(med), _p(sh), _p(pred)))
        return {"median": med, "sharpened": sh, "predicate": pred}

    def region_grow(self, img, seeds, params=None, gpu=0):
        img = np.ascontiguousarray(img, np.float32)
        h, w = img.shape
        params = params or default_params()
        seeds = np.ascontiguousarray(seeds, np.int32).reshape(-1, 2)
        out = np.empty((h, w), np.uint8)
        it = C.c_int(0)
        rc = lib().nm03_region_grow(self._h, gpu, _p(img), w, h, C.byref(params), _p(seeds), len(seeds), _p(out),
                                    C.cast(C.byref(it), C.c_void_p))
        _check(rc)
        return out, it.value

    def morphology(self, mask, size=3, se_shape=0, op=0, gpu=0):
        mask = np.ascontiguousarray(mask, np.uint8)
        h, w = mask.shape
        out = np.empty((h, w), np.uint8)
        _check(lib().nm03_morphology(self._h, gpu, _p(mask), w, h, size, se_shape, op, _p(out)))
        return out


def shard_range(n_items, world, rank):
    """Contiguous patient-range sharding: GPU g of G gets items [floor(g*P/G), floor((g+1)*P/G)).
    The reference's analogue is the sorted patient list walked by one process (main_parallel.cpp:248,373)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    return (rank * n_items) // world, ((rank + 1) * n_items) // world


def unpack_bits(bits, w):
    """bit-packed masks (..., h, ceil(w/32)) uint32 -> (..., h, w) uint8 {0,1} (nm03_outputs.dilated_bits)."""
    b = np.unpackbits(np.ascontiguousarray(bits).view(np.uint8).reshape(bits.shape[:-1] + (-1,)), axis=-1, bitorder="little")
    return b[..., :w]


def unpack_predicate(pred, w):
    """bit-plane (h, ceil(w/32)) uint32 -> (h, w) uint8."""
    bits = np.unpackbits(pred.view(np.uint8).reshape(pred.shape[0], -1), axis=1, bitorder="little")
    return bits[:, :w]
