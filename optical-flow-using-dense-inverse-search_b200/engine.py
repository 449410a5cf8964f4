"""ctypes binding of the C ABI in include/dis_b200.h (libdis_b200.so, hand-written sm_100a kernels).

This is the Python face of the product path: it only marshals pointers.  There is no CPU
implementation behind it — if the library is missing or no B200 is visible every call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DIS_B200_LIB") or os.path.join(_HERE, "libdis_b200.so")   # the override is for tuning builds (tools/variants.sh)

_u8p = C.POINTER(C.c_uint8)
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


class DisParams(C.Structure):
    """dis_params — the parameters of /root/reference/src/main.cpp:63-72."""
    _fields_ = [("max_iter", C.c_int), ("patch_size", C.c_int), ("coarsest_scale", C.c_int),
                ("finest_scale", C.c_int), ("patch_overlap", C.c_float), ("patch_norm", C.c_int),
                ("conv_eps", C.c_float), ("steps", C.c_int), ("outlier_thresh", C.c_float),
                ("reserved", C.c_int * 3)]


class DisLevelInfo(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("scale", "width", "height", "tmp_w", "tmp_h", "npw", "nph", "offw", "offh",
                                       "steps", "num_patches")]


class DisStats(C.Structure):
    _fields_ = [(n, C.c_ulonglong) for n in ("patches", "patch_iterations", "sampled_iterations", "resets",
                                             "start_oob", "kernel_launches", "sampled_converge", "sampled_walk")]


class DisError(RuntimeError):
    pass


_lib = None

# every symbol include/dis_b200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "dis_default_params", "dis_last_error", "dis_version", "dis_tune", "dis_create", "dis_destroy", "dis_run_pair_u8",
    "dis_run_batch_u8", "dis_run_sequence_u8", "dis_run_pyramids", "dis_construct_pyramid", "dis_upload_frames_u8",
    "dis_run_resident", "dis_download_flows", "dis_synchronize", "dis_timer_start", "dis_timer_stop_ms",
    "dis_last_search_ms", "dis_last_search_level_ms", "dis_last_search_phase_ms", "dis_num_levels", "dis_get_level_info", "dis_get_patches", "dis_get_level_flow",
    "dis_get_plane", "dis_get_stats", "dis_grid_create", "dis_grid_create_single", "dis_grid_destroy", "dis_grid_init", "dis_grid_set_target",
    "dis_grid_init_from_prev_flow", "dis_grid_set_patch_init", "dis_grid_inverse_search", "dis_grid_densify",
    "dis_grid_num_patches", "dis_grid_get_patch", "dis_write_flo", "dis_read_flo", "dis_read_flo_size",
    "dis_color_code", "dis_color_code_pair", "dis_tiled_begin", "dis_tiled_search", "dis_tiled_get_u", "dis_tiled_set_u",
    "dis_tiled_densify", "dis_tiled_finish",
    "dis_pool_create", "dis_pool_destroy", "dis_pool_size", "dis_pool_device", "dis_pool_last_error", "dis_pool_shard",
    "dis_pool_run_batch_u8", "dis_pool_run_sequence_u8", "dis_pool_get_stats", "dis_pool_last_ms", "dis_pool_host_alloc",
    "dis_pool_host_free", "dis_pool_run_pair_tiled_u8", "dis_tiled_fetch_u_rows", "dis_tiled_densify_rows",
    "dis_tiled_finish_rows", "dis_tiled_end", "dis_tiled_last_ms", "dis_pool_last_tiled_device_ms",
    "dis_color_code_device_noalloc", "dis_set_host_slices", "dis_pool_submit_sequence_u8", "dis_pool_submit_batch_u8",
    "dis_pool_wait", "dis_pool_wait_all",
]


def load_library() -> C.CDLL:
    """Load libdis_b200.so; raise loudly when it was not built (no fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DisError(f"{LIB_PATH} not built: run __graft_entry__.build() (make -C csrc); there is no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        lib.dis_last_error.restype = C.c_char_p
        lib.dis_version.restype = C.c_char_p
        lib.dis_destroy.restype = None
        lib.dis_grid_destroy.restype = None
        lib.dis_destroy.argtypes = [C.c_void_p]
        lib.dis_grid_destroy.argtypes = [C.c_void_p]
        _lib = lib
    return _lib


def _chk(rc: int) -> None:
    if rc != 0:
        raise DisError(f"dis error {rc}: {load_library().dis_last_error().decode()}")


def default_params(**kw) -> DisParams:
    p = DisParams()
    load_library().dis_default_params(C.byref(p))
    for k, v in kw.items():
        if not hasattr(p, k):
            raise TypeError(f"unknown parameter {k}")
        setattr(p, k, v)
    return p


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


class Engine:
    """One dis_engine handle (one CUDA stream, one device)."""

    def __init__(self, device: int = 0, **params):
        self.lib = load_library()
        self.params = default_params(**params)
        self._h = C.c_void_p()
        _chk(self.lib.dis_create(C.byref(self.params), int(device), C.byref(self._h)))
        self._shape = None
        self.device = int(device)

    def close(self):
        if self._h:
            self.lib.dis_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- whole path, host buffers ---------------------------------------------------------
    def run_pair(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        assert a.ndim == 2 and a.shape == b.shape
        h, w = a.shape
        out = np.empty((h, w, 2), np.float32)
        _chk(self.lib.dis_run_pair_u8(self._h, a.ctypes.data_as(_u8p), b.ctypes.data_as(_u8p), w, h, w,
                                      out.ctypes.data_as(_fp)))
        return out

    # ---- one large frame tiled by grid rows over several engines ------------
    def tiled_begin(self, a: np.ndarray, b: np.ndarray) -> None:
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        h, w = a.shape
        _chk(self.lib.dis_tiled_begin(self._h, a.ctypes.data_as(_u8p), b.ctypes.data_as(_u8p), w, h, w))
        self._shape = (1, h, w)

    def tiled_search(self, level: int, row0: int, row1: int) -> None:
        _chk(self.lib.dis_tiled_search(self._h, int(level), int(row0), int(row1)))

    def tiled_get_u(self, level: int, device_ptr: int) -> None:
        _chk(self.lib.dis_tiled_get_u(self._h, int(level), C.c_void_p(device_ptr)))

    def tiled_set_u(self, level: int, device_ptr: int) -> None:
        _chk(self.lib.dis_tiled_set_u(self._h, int(level), C.c_void_p(device_ptr)))

    def tiled_densify(self, level: int) -> None:
        _chk(self.lib.dis_tiled_densify(self._h, int(level)))

    def tiled_finish(self) -> np.ndarray:
        _, h, w = self._shape
        out = np.empty((h, w, 2), np.float32)
        _chk(self.lib.dis_tiled_finish(self._h, out.ctypes.data_as(_fp)))
        return out

    def color_code(self, flow: np.ndarray, maxmotion: float = -1.0) -> np.ndarray:
        """draw_optical_flow (/root/reference/src/color_coding.cpp:83-117): (h,w,2) fp32 -> (h,w,3) u8 B,G,R."""
        flow = np.ascontiguousarray(flow, np.float32)
        h, w, _ = flow.shape
        out = np.empty((h, w, 3), np.uint8)
        _chk(self.lib.dis_color_code(self._h, flow.ctypes.data_as(_fp), w, h, C.c_float(maxmotion), out.ctypes.data_as(_u8p)))
        return out

    def color_code_pair(self, pair: int, h: int, w: int, maxmotion: float = -1.0) -> np.ndarray:
        """Colour coding of the device-resident flow of pair `pair` of the last run."""
        out = np.empty((h, w, 3), np.uint8)
        _chk(self.lib.dis_color_code_pair(self._h, pair, C.c_float(maxmotion), out.ctypes.data_as(_u8p)))
        return out

    def set_host_slices(self, enabled: bool) -> None:
        """dis_set_host_slices: host-buffer calls return the flows in slices (default) or in one copy at the end."""
        _chk(self.lib.dis_set_host_slices(self._h, C.c_int(1 if enabled else 0)))

    def run_batch(self, a: np.ndarray, b: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        n, h, w = a.shape
        if out is None:
            out = np.empty((n, h, w, 2), np.float32)
        pa = (C.c_void_p * n)(*[a.ctypes.data + i * h * w for i in range(n)])
        pb = (C.c_void_p * n)(*[b.ctypes.data + i * h * w for i in range(n)])
        _chk(self.lib.dis_run_batch_u8(self._h, pa, pb, n, w, h, w, out.ctypes.data_as(_fp)))
        return out

    def run_sequence(self, frames: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape
        if out is None:
            out = np.empty((n - 1, h, w, 2), np.float32)
        _chk(self.lib.dis_run_sequence_u8(self._h, frames.ctypes.data_as(_u8p), n, w, h, w, C.c_size_t(h * w),
                                          out.ctypes.data_as(_fp)))
        return out

    def run_sequence_ptr(self, frames_ptr: int, n: int, w: int, h: int, out_ptr: int) -> None:
        """Same as run_sequence on raw (e.g. pinned) host addresses."""
        _chk(self.lib.dis_run_sequence_u8(self._h, C.cast(C.c_void_p(frames_ptr), _u8p), n, w, h, w, C.c_size_t(h * w),
                                          C.cast(C.c_void_p(out_ptr), _fp)))

    # ---- the OpticalFlowClass contract ------------------------------------------------------
    def run_pyramids(self, pyr_first, pyr_second, img_padding: int, width: int, height: int) -> np.ndarray:
        """pyr_* = (img, dx, dy) lists indexed by absolute scale (padded fp32 planes)."""
        def arr(planes):
            keep = [None if p is None else _f32(p) for p in planes]
            ptrs = (C.c_void_p * len(keep))(*[None if p is None else p.ctypes.data for p in keep])
            return keep, ptrs
        keep, six = [], []
        for planes in (*pyr_first, *pyr_second):
            k, p = arr(planes)
            keep.append(k); six.append(p)
        f = self.params.finest_scale
        out = np.empty((height >> f, width >> f, 2), np.float32)
        _chk(self.lib.dis_run_pyramids(self._h, *six, int(img_padding), out.ctypes.data_as(_fp), int(width), int(height)))
        return out

    def construct_pyramid(self, img: np.ndarray, lv_f: int, img_padding: int):
        img = _f32(img)
        H, W = img.shape
        mk = lambda: [np.empty(((H >> i) + 2 * img_padding, (W >> i) + 2 * img_padding), np.float32) for i in range(lv_f + 1)]
        P, Px, Py = mk(), mk(), mk()
        ptr = lambda L: (C.c_void_p * len(L))(*[x.ctypes.data for x in L])
        _chk(self.lib.dis_construct_pyramid(self._h, img.ctypes.data_as(_fp), W, H, int(lv_f), int(img_padding),
                                            ptr(P), ptr(Px), ptr(Py)))
        return P, Px, Py

    # ---- device-resident pipeline -------------------------------------------------------------
    def upload_frames(self, frames: np.ndarray, first, second) -> None:
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape
        first = np.ascontiguousarray(first, np.int32); second = np.ascontiguousarray(second, np.int32)
        _chk(self.lib.dis_upload_frames_u8(self._h, frames.ctypes.data_as(_u8p), n, w, h, w, C.c_size_t(h * w),
                                           first.ctypes.data_as(_ip), second.ctypes.data_as(_ip), len(first)))
        self._shape = (len(first), h, w)

    def run_resident(self) -> None:
        _chk(self.lib.dis_run_resident(self._h))

    def download_flows(self, out: np.ndarray | None = None) -> np.ndarray:
        n, h, w = self._shape
        if out is None:
            out = np.empty((n, h, w, 2), np.float32)
        _chk(self.lib.dis_download_flows(self._h, out.ctypes.data_as(_fp)))
        return out

    def synchronize(self) -> None:
        _chk(self.lib.dis_synchronize(self._h))

    def timer_start(self) -> None:
        _chk(self.lib.dis_timer_start(self._h))

    def timer_stop_ms(self) -> float:
        ms = C.c_float()
        _chk(self.lib.dis_timer_stop_ms(self._h, C.byref(ms)))
        return ms.value

    def last_search_ms(self):
        ms, n = C.c_float(), C.c_int()
        _chk(self.lib.dis_last_search_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def last_search_phase_ms(self, level: int, phase: int) -> float:
        """Device time of one kernel of a level: 0 k_converge, 1 k_inverse_search (walk), 2 k_stragglers."""
        ms = C.c_float()
        _chk(self.lib.dis_last_search_phase_ms(self._h, int(level), int(phase), C.byref(ms)))
        return ms.value

    def last_search_level_ms(self, level: int) -> float:
        ms = C.c_float()
        _chk(self.lib.dis_last_search_level_ms(self._h, int(level), C.byref(ms)))
        return ms.value

    # ---- introspection ------------------------------------------------------------------------
    def num_levels(self) -> int:
        return self.lib.dis_num_levels(self._h)

    def level_info(self, level: int) -> dict:
        li = DisLevelInfo()
        _chk(self.lib.dis_get_level_info(self._h, level, C.byref(li)))
        return {n: getattr(li, n) for n, _ in DisLevelInfo._fields_}

    def patches(self, pair: int, level: int):
        n = self.level_info(level)["num_patches"]
        ref = np.empty((n, 2), np.float32); u = np.empty((n, 2), np.float32); it = np.empty(n, np.int32)
        _chk(self.lib.dis_get_patches(self._h, pair, level, ref.ctypes.data_as(_fp), u.ctypes.data_as(_fp),
                                      it.ctypes.data_as(_ip)))
        return ref, u, it

    def level_flow(self, pair: int, level: int) -> np.ndarray:
        li = self.level_info(level)
        out = np.empty((li["height"], li["width"], 2), np.float32)
        _chk(self.lib.dis_get_level_flow(self._h, pair, level, out.ctypes.data_as(_fp)))
        return out

    def plane(self, frame: int, scale: int, which: int, shape) -> np.ndarray:
        out = np.empty(shape, np.float32)
        _chk(self.lib.dis_get_plane(self._h, frame, scale, which, out.ctypes.data_as(_fp)))
        return out

    def stats(self) -> dict:
        s = DisStats()
        _chk(self.lib.dis_get_stats(self._h, C.byref(s)))
        return {n: int(getattr(s, n)) for n, _ in DisStats._fields_}


class Pool:
    """Several GPUs behind one handle (dis_pool_*): one engine + host thread per device, pairs sharded contiguously,
    no collective (/root/reference/src/main.cpp:102 — the loop body shares no state between pairs)."""

    def __init__(self, devices, **params):
        self.lib = load_library()
        self.lib.dis_pool_last_error.restype = C.c_char_p
        self.lib.dis_pool_destroy.restype = None
        self.lib.dis_pool_destroy.argtypes = [C.c_void_p]
        self.params = default_params(**params)
        self.devices = [int(d) for d in devices]
        self._h = C.c_void_p()
        dv = (C.c_int * len(self.devices))(*self.devices)
        self._chk(self.lib.dis_pool_create(C.byref(self.params), dv, len(self.devices), C.byref(self._h)))

    def _chk(self, rc):
        if rc != 0:
            raise DisError(f"dis_b200 pool error {rc}: {self.lib.dis_pool_last_error().decode()}")

    def close(self):
        if self._h:
            self.lib.dis_pool_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def run_sequence(self, frames: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape
        if out is None:
            out = np.empty((n - 1, h, w, 2), np.float32)
        self._chk(self.lib.dis_pool_run_sequence_u8(self._h, frames.ctypes.data_as(_u8p), n, w, h, w, C.c_size_t(h * w),
                                                    out.ctypes.data_as(_fp)))
        return out

    def run_sequence_ptr(self, frames_ptr: int, n: int, w: int, h: int, out_ptr: int) -> None:
        self._chk(self.lib.dis_pool_run_sequence_u8(self._h, C.cast(C.c_void_p(frames_ptr), _u8p), n, w, h, w, C.c_size_t(h * w),
                                                    C.cast(C.c_void_p(out_ptr), _fp)))

    def run_batch(self, a: np.ndarray, b: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        n, h, w = a.shape
        if out is None:
            out = np.empty((n, h, w, 2), np.float32)
        pa = (C.c_void_p * n)(*[a.ctypes.data + i * h * w for i in range(n)])
        pb = (C.c_void_p * n)(*[b.ctypes.data + i * h * w for i in range(n)])
        self._chk(self.lib.dis_pool_run_batch_u8(self._h, pa, pb, n, w, h, w, out.ctypes.data_as(_fp)))
        return out

    # -- asynchronous form: a whole call per engine, tickets (dis_pool_submit_* / dis_pool_wait) --
    def submit_sequence_ptr(self, frames_ptr: int, n: int, w: int, h: int, out_ptr: int) -> int:
        """Queue one whole sequence call on the idlest engine; the buffers belong to the pool until wait(ticket)."""
        t = C.c_int(0)
        self._chk(self.lib.dis_pool_submit_sequence_u8(self._h, C.cast(C.c_void_p(frames_ptr), _u8p), n, w, h, w, C.c_size_t(h * w),
                                                       C.cast(C.c_void_p(out_ptr), _fp), C.byref(t)))
        return t.value

    def submit_sequence(self, frames: np.ndarray, out: np.ndarray) -> int:
        assert frames.dtype == np.uint8 and frames.flags.c_contiguous and out.dtype == np.float32 and out.flags.c_contiguous
        n, h, w = frames.shape
        assert out.shape == (n - 1, h, w, 2)
        return self.submit_sequence_ptr(frames.ctypes.data, n, w, h, out.ctypes.data)

    def submit_batch(self, a: np.ndarray, b: np.ndarray, out: np.ndarray) -> int:
        assert a.dtype == np.uint8 and a.flags.c_contiguous and b.dtype == np.uint8 and b.flags.c_contiguous
        n, h, w = a.shape
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.shape == (n, h, w, 2)
        pa = (C.c_void_p * n)(*[a.ctypes.data + i * h * w for i in range(n)])
        pb = (C.c_void_p * n)(*[b.ctypes.data + i * h * w for i in range(n)])
        t = C.c_int(0)
        self._chk(self.lib.dis_pool_submit_batch_u8(self._h, pa, pb, n, w, h, w, out.ctypes.data_as(_fp), C.byref(t)))
        return t.value

    def wait(self, ticket: int) -> dict:
        """Wait for one submission; returns its stats plus the worker that ran it."""
        st = DisStats()
        wk = C.c_int(-1)
        self._chk(self.lib.dis_pool_wait(self._h, C.c_int(ticket), C.byref(st), C.byref(wk)))
        d = {k: int(getattr(st, k)) for k, _ in DisStats._fields_}
        d["worker"] = wk.value
        return d

    def wait_all(self) -> None:
        self._chk(self.lib.dis_pool_wait_all(self._h))

    def run_pair_tiled(self, a: np.ndarray, b: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """One large pair over all the pool's GPUs: bands of grid rows + per-level halo exchange over peer memory."""
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        h, w = a.shape
        if out is None:
            out = np.empty((h, w, 2), np.float32)
        self._chk(self.lib.dis_pool_run_pair_tiled_u8(self._h, a.ctypes.data_as(_u8p), b.ctypes.data_as(_u8p), w, h, w,
                                                      out.ctypes.data_as(_fp)))
        return out

    def last_tiled_device_ms(self) -> float:
        ms = C.c_double()
        self._chk(self.lib.dis_pool_last_tiled_device_ms(self._h, C.byref(ms)))
        return float(ms.value)

    def run_pair_tiled_ptr(self, a_ptr: int, b_ptr: int, w: int, h: int, out_ptr: int) -> None:
        self._chk(self.lib.dis_pool_run_pair_tiled_u8(self._h, C.cast(C.c_void_p(a_ptr), _u8p), C.cast(C.c_void_p(b_ptr), _u8p),
                                                      w, h, w, C.cast(C.c_void_p(out_ptr), _fp)))

    def run_batch_ptr(self, a_ptr: int, b_ptr: int, n: int, w: int, h: int, out_ptr: int) -> None:
        """n pairs stored back to back at a_ptr / b_ptr (e.g. pinned host memory)."""
        pa = (C.c_void_p * n)(*[a_ptr + i * h * w for i in range(n)])
        pb = (C.c_void_p * n)(*[b_ptr + i * h * w for i in range(n)])
        self._chk(self.lib.dis_pool_run_batch_u8(self._h, pa, pb, n, w, h, w, C.cast(C.c_void_p(out_ptr), _fp)))

    def stats(self) -> dict:
        s = DisStats()
        self._chk(self.lib.dis_pool_get_stats(self._h, C.byref(s)))
        return {n: int(getattr(s, n)) for n, _ in DisStats._fields_}

    def last_ms(self):
        wall = C.c_double()
        per = (C.c_double * len(self.devices))()
        self._chk(self.lib.dis_pool_last_ms(self._h, C.byref(wall), per))
        return float(wall.value), [float(x) for x in per]

    def workers(self):
        out = []
        for i in range(len(self.devices)):
            d, n = C.c_int(), C.c_int()
            self._chk(self.lib.dis_pool_device(self._h, i, C.byref(d), C.byref(n)))
            out.append({"device": int(d.value), "numa_cpus": int(n.value)})
        return out

    def host_alloc(self, worker: int, nbytes: int) -> int:
        p = C.c_void_p()
        self._chk(self.lib.dis_pool_host_alloc(self._h, int(worker), C.c_size_t(nbytes), C.byref(p)))
        return int(p.value)

    def host_free(self, ptr: int) -> None:
        self._chk(self.lib.dis_pool_host_free(self._h, C.c_void_p(ptr)))


NOALLOC_MAX_PIXELS = 8192 * 4096   # DIS_NOALLOC_MAX_PIXELS of include/dis_b200.h


def color_code_device_noalloc(device: int, d_flow_ptr: int, w: int, h: int, d_bgr_ptr: int, maxmotion: float = -1.0,
                              stream: int = 0) -> None:
    """dis_color_code_device_noalloc: colour coding of a DEVICE flow field into a DEVICE image, enqueued on `stream`, with
    no allocation of any kind (fixed scratch inside the library; w*h <= NOALLOC_MAX_PIXELS).  No counterpart in the
    reference; same bytes as Engine.color_code.  Calls overlapping on one GPU race on the shared scratch."""
    lib = load_library()
    _chk(lib.dis_color_code_device_noalloc(C.c_int(device), C.c_void_p(d_flow_ptr), C.c_int(w), C.c_int(h), C.c_float(maxmotion),
                                           C.c_void_p(d_bgr_ptr), C.c_void_p(stream)))


def write_flo(path: str, flow: np.ndarray) -> None:
    flow = _f32(flow)
    _chk(load_library().dis_write_flo(path.encode(), flow.ctypes.data_as(_fp), flow.shape[1], flow.shape[0]))


def read_flo(path: str) -> np.ndarray:
    lib = load_library()
    w, h = C.c_int(), C.c_int()
    _chk(lib.dis_read_flo_size(path.encode(), C.byref(w), C.byref(h)))
    out = np.empty((h.value, w.value, 2), np.float32)
    _chk(lib.dis_read_flo(path.encode(), out.ctypes.data_as(_fp), w.value, h.value))
    return out
