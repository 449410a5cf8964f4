"""ctypes bindings for oracle/libdis_oracle.so (C restatement, "port") and
oracle/_ref/libdis_ref.so (the reference's own engine sources, "reference").
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PORT = os.path.join(_HERE, "libdis_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libdis_ref.so")
_REFVIZ = os.path.join(_HERE, "_ref", "libdis_refviz.so")
_fpp = C.POINTER(C.POINTER(C.c_float))
_fp = C.POINTER(C.c_float)


def build(ref: bool = True) -> None:
    """Compile the C restatement and, where /root/reference exists, oracle/_ref."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle"] + (["ref"] if ref else []))


def load_port() -> C.CDLL:
    if not os.path.exists(_PORT):
        build(ref=False)
    return C.CDLL(_PORT)


def load_ref():
    """The compiled reference, or None when it was never built (no /root/reference here)."""
    if not os.path.exists(_REF):
        if os.path.isdir("/root/reference/src"):
            build(ref=True)
        else:
            return None
    lib = C.CDLL(_REF)
    assert lib.ref_new_is_zeroing() == 1, "zero-filling operator new[] not bound inside libdis_ref.so"
    return lib


def load_refviz():
    """The reference's own colour coding + .flo I/O (src/color_coding.cpp, src/IO_flow.cpp compiled unmodified), or None."""
    if not os.path.exists(_REFVIZ):
        if os.path.isdir("/root/reference/src"):
            build(ref=True)
        else:
            return None
    return C.CDLL(_REFVIZ)


class RefViz:
    """draw_optical_flow / SaveFlowFile / ReadFlowFile of the reference itself."""

    def __init__(self):
        self.lib = load_refviz()
        if self.lib is None:
            raise RuntimeError("oracle/_ref/libdis_refviz.so not available")

    def color_code(self, flow, maxmotion=-1.0):
        flow = np.ascontiguousarray(flow, np.float32)
        h, w, _ = flow.shape
        out = np.empty((h, w, 3), np.uint8)
        self.lib.ref_color_code(flow.ctypes.data_as(_fp), w, h, C.c_float(maxmotion), out.ctypes.data_as(C.c_void_p))
        return out

    def save_flo(self, path, flow):
        flow = np.ascontiguousarray(flow, np.float32)
        self.lib.ref_save_flo(flow.ctypes.data_as(_fp), flow.shape[1], flow.shape[0], str(path).encode())

    def read_flo(self, path, w, h):
        out = np.zeros((h, w, 2), np.float32)
        self.lib.ref_read_flo(str(path).encode(), out.ctypes.data_as(_fp), w, h)
        return out


def _ptrs(arrs):
    a = (_fp * len(arrs))()
    for i, x in enumerate(arrs):
        a[i] = x.ctypes.data_as(_fp) if x is not None else None
    return a


class Oracle:
    """kind='port' (dis_oracle.c) or 'reference' (reference sources via oracle/_ref)."""

    def __init__(self, kind: str = "port"):
        self.kind = kind
        self.port = load_port()
        self.ref = load_ref() if kind == "reference" else None
        if kind == "reference" and self.ref is None:
            raise RuntimeError("oracle/_ref/libdis_ref.so not available")
        self.port.oracle_steps.argtypes = [C.c_int, C.c_float]

    # ---- pre-processing (always the C restatement: the reference's is OpenCV calls) ----
    def div_pad(self, dim, coarsest):
        return self.port.oracle_div_pad(int(dim), int(coarsest))

    def pad_convert(self, img, coarsest):
        h, w = img.shape
        img = np.ascontiguousarray(img, np.uint8)
        pw, ph = self.div_pad(w, coarsest), self.div_pad(h, coarsest)
        out = np.empty((h + ph, w + pw), np.float32)
        self.port.oracle_pad_convert(img.ctypes.data_as(C.c_void_p), w, h, w, pw, ph, out.ctypes.data_as(_fp))
        return out, pw, ph

    def sobel(self, p):
        p = np.ascontiguousarray(p, np.float32)
        gx, gy = np.empty_like(p), np.empty_like(p)
        self.port.oracle_sobel(p.ctypes.data_as(_fp), p.shape[1], p.shape[0], gx.ctypes.data_as(_fp), gy.ctypes.data_as(_fp))
        return gx, gy

    def pyramid(self, fimg, levels, pad):
        """fimg: float32 HxW (divisibility padded). Returns (P, Px, Py) lists of padded planes."""
        fimg = np.ascontiguousarray(fimg, np.float32)
        H, W = fimg.shape
        mk = lambda: [np.empty(((H >> i) + 2 * pad, (W >> i) + 2 * pad), np.float32) for i in range(levels)]
        P, Px, Py = mk(), mk(), mk()
        self.port.oracle_construct_pyramid(fimg.ctypes.data_as(_fp), W, H, levels, pad, _ptrs(P), _ptrs(Px), _ptrs(Py))
        return P, Px, Py

    def level_info(self, W, H, scale, pad, ps, overlap):
        out = (C.c_int * 8)()
        self.port.oracle_level_info.argtypes = [C.c_int] * 5 + [C.c_float, C.POINTER(C.c_int)]
        self.port.oracle_level_info(W, H, scale, pad, ps, overlap, out)
        return dict(zip(("w", "h", "npw", "nph", "offw", "offh", "steps", "tmp_w"), list(out)))

    # ---- engine ----
    def optical_flow(self, pyrA, pyrB, pad, W, H, coarsest=3, finest=0, iterations=1000, patch_size=8,
                     overlap=0.7, patnorm=True, dump=False):
        """pyrA/pyrB = (P, Px, Py) lists indexed by absolute scale.  Returns flow (h,w,2) and, with
        dump=True, a dict of per-level arrays keyed by level index s = scale - finest."""
        nl = coarsest - finest + 1
        h, w = H >> finest, W >> finest
        out = np.zeros((h, w, 2), np.float32)
        six = [_ptrs(x) for x in (*pyrA, *pyrB)]
        info = [self.level_info(W, H, finest + s, pad, patch_size, overlap) for s in range(nl)]
        d = None
        if dump:
            d = {"flow": [np.zeros((i["h"], i["w"], 2), np.float32) for i in info],
                 "ref": [np.zeros((i["npw"] * i["nph"], 2), np.float32) for i in info],
                 "info": info}
        if self.kind == "port":
            f = self.port.oracle_optical_flow
            f.argtypes = [_fpp] * 6 + [C.c_int, _fp] + [C.c_int] * 6 + [C.c_float, C.c_int, _fpp, _fpp, _fpp, C.c_void_p]
            if dump:
                d["u"] = [np.zeros_like(r) for r in d["ref"]]
                d["iters"] = [np.zeros(len(r), np.int32) for r in d["ref"]]
                it = (C.POINTER(C.c_int) * nl)(*[x.ctypes.data_as(C.POINTER(C.c_int)) for x in d["iters"]])
                f(*six, pad, out.ctypes.data_as(_fp), W, H, coarsest, finest, iterations, patch_size, overlap,
                  int(patnorm), _ptrs(d["flow"]), _ptrs(d["ref"]), _ptrs(d["u"]), C.cast(it, C.c_void_p))
            else:
                f(*six, pad, out.ctypes.data_as(_fp), W, H, coarsest, finest, iterations, patch_size, overlap,
                  int(patnorm), None, None, None, None)
        else:
            if dump:
                f = self.ref.ref_optical_flow_levels
                f.argtypes = [_fpp] * 6 + [C.c_int, _fp] + [C.c_int] * 6 + [C.c_float, C.c_int, _fpp, _fpp, _fpp, C.c_void_p]
                d["query"] = [np.zeros_like(r) for r in d["ref"]]
                f(*six, pad, out.ctypes.data_as(_fp), W, H, coarsest, finest, iterations, patch_size, overlap,
                  int(patnorm), _ptrs(d["flow"]), _ptrs(d["ref"]), _ptrs(d["query"]), None)
            else:
                f = self.ref.ref_optical_flow
                f.argtypes = [_fpp] * 6 + [C.c_int, _fp] + [C.c_int] * 6 + [C.c_float, C.c_int]
                f(*six, pad, out.ctypes.data_as(_fp), W, H, coarsest, finest, iterations, patch_size, overlap, int(patnorm))
        return (out, d) if dump else out

    def color_code(self, flow, maxmotion=-1.0):
        """draw_optical_flow (src/color_coding.cpp:83-117) -> (h,w,3) u8 B,G,R."""
        flow = np.ascontiguousarray(flow, np.float32)
        h, w, _ = flow.shape
        out = np.empty((h, w, 3), np.uint8)
        f = self.port.oracle_color_code
        f.argtypes = [_fp, C.c_int, C.c_int, C.c_float, C.c_void_p]
        f(flow.ctypes.data_as(_fp), w, h, maxmotion, out.ctypes.data)
        return out

    def run_pair_u8(self, a, b, iterations=1000, patch_size=8, coarsest=3, finest=0, overlap=0.7, patnorm=True,
                    dump=False):
        """Whole path of src/main.cpp:135-198 (finest == 0): pad, pyramids, engine, crop."""
        fa, pw, ph = self.pad_convert(a, coarsest)
        fb, _, _ = self.pad_convert(b, coarsest)
        H, W = fa.shape
        pa = self.pyramid(fa, coarsest + 1, patch_size)
        pb = self.pyramid(fb, coarsest + 1, patch_size)
        r = self.optical_flow(pa, pb, patch_size, W, H, coarsest, finest, iterations, patch_size, overlap, patnorm, dump)
        flow, d = r if dump else (r, None)
        sc = 1 << finest
        top, left = (ph // 2) // sc, (pw // 2) // sc
        crop = flow[top:top + a.shape[0] // sc, left:left + a.shape[1] // sc]
        return (np.ascontiguousarray(crop), d) if dump else np.ascontiguousarray(crop)
