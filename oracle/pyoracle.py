"""ctypes bindings for the CPU oracles -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Two libraries with the same C surface:

  * oracle/_ref/liborbref.so      prefix `orbref_`  the reference's ORBextractor.cc, verbatim
  * oracle/_build/liborb_oracle.so prefix `orbo_`   our own restatement ("port")
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "liborbref.so")
PORT_SO = os.path.join(HERE, "_build", "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build(target: str = "all") -> None:
    subprocess.run(["make", "-s", "-f", os.path.join(HERE, "Makefile"), target], check=True)


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def have_port() -> bool:
    return os.path.exists(PORT_SO)


_libs = {}


def _lib(kind: str):
    if kind in _libs:
        return _libs[kind]
    path, pre = (REF_SO, "orbref_") if kind == "ref" else (PORT_SO, "orbo_")
    lib = C.CDLL(path)
    f = lambda n: getattr(lib, pre + n)
    vp, i, u8p, ip, fp, sz = C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t
    f("create").restype = vp
    f("create").argtypes = [i, C.c_float, i, i, i]
    f("destroy").restype = None
    f("destroy").argtypes = [vp]
    f("set_malloc_mode").restype = None
    f("set_malloc_mode").argtypes = [i]
    f("tables").restype = i
    f("tables").argtypes = [vp, fp, fp, fp, fp, ip]
    f("umax").restype = i
    f("umax").argtypes = [vp, ip]
    f("pattern").restype = i
    f("pattern").argtypes = [vp, ip]
    f("extract").restype = i
    f("extract").argtypes = [vp, u8p, i, i, sz, i, i, vp, u8p, i, ip]
    f("level_size").restype = i
    f("level_size").argtypes = [vp, i, ip, ip]
    f("get_level").restype = i
    f("get_level").argtypes = [vp, i, i, u8p, sz]
    f("stages").restype = i
    f("stages").argtypes = [vp, u8p, i, i, sz]
    f("stage_candidates").restype = i
    f("stage_candidates").argtypes = [vp, i, ip, i]
    f("stage_survivors").restype = i
    f("stage_survivors").argtypes = [vp, i, vp, i]
    _libs[kind] = (lib, pre)
    return _libs[kind]


class OracleExtractor:
    """Mirror of ORB_SLAM3::ORBextractor (include/ORBextractor.h:43-109) on the CPU oracle."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 kind="ref"):
        self.kind = kind
        self.lib, self.pre = _lib(kind)
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = self._f("create")(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        self.cap = nfeatures + 64 * nlevels + 1024

    def _f(self, n):
        return getattr(self.lib, self.pre + n)

    def close(self):
        if self.h:
            self._f("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_malloc_mode(self, on: bool):
        self._f("set_malloc_mode")(1 if on else 0)

    def tables(self):
        n = self.nlevels
        a, b, c, d = (np.zeros(n, np.float32) for _ in range(4))
        e = np.zeros(n, np.int32)
        self._f("tables")(self.h, a.ctypes.data, b.ctypes.data, c.ctypes.data, d.ctypes.data,
                          e.ctypes.data)
        return dict(scale=a, inv_scale=b, sigma2=c, inv_sigma2=d, n_per_level=e)

    def umax(self):
        u = np.zeros(16, np.int32)
        self._f("umax")(self.h, u.ctypes.data)
        return u

    def pattern(self):
        p = np.zeros(1024, np.int32)
        self._f("pattern")(self.h, p.ctypes.data)
        return p.reshape(512, 2)

    def __call__(self, img: np.ndarray | None, lapping=(0, 0)):
        """Returns (mono_index, keypoints[KP_DTYPE], descriptors[n,32] u8)."""
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        if img is None or img.size == 0:
            mono = self._f("extract")(self.h, None, 0, 0, 0, lapping[0], lapping[1],
                                      kps.ctypes.data, desc.ctypes.data, self.cap, C.byref(n))
        else:
            assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
            mono = self._f("extract")(self.h, img.ctypes.data, img.shape[1], img.shape[0],
                                      img.strides[0], lapping[0], lapping[1], kps.ctypes.data,
                                      desc.ctypes.data, self.cap, C.byref(n))
        assert n.value >= 0, "oracle capacity too small"
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def level(self, lvl: int, pad: int = 0) -> np.ndarray:
        w, h = C.c_int(), C.c_int()
        self._f("level_size")(self.h, lvl, C.byref(w), C.byref(h))
        out = np.zeros((h.value + 2 * pad, w.value + 2 * pad), np.uint8)
        self._f("get_level")(self.h, lvl, pad, out.ctypes.data, out.strides[0])
        return out

    def stages(self, img: np.ndarray):
        """Returns (candidates[level] -> int32[n,3] (x,y,response) relative to (16,16),
        survivors[level] -> KP_DTYPE[n] in level coordinates, with angle)."""
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        r = self._f("stages")(self.h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0])
        assert r == 0
        cands, survs = [], []
        for l in range(self.nlevels):
            n = self._f("stage_candidates")(self.h, l, None, 0)
            c = np.zeros((n, 3), np.int32)
            if n:
                self._f("stage_candidates")(self.h, l, c.ctypes.data, n)
            cands.append(c)
            n = self._f("stage_survivors")(self.h, l, None, 0)
            s = np.zeros(n, KP_DTYPE)
            if n:
                self._f("stage_survivors")(self.h, l, s.ctypes.data, n)
            survs.append(s)
        return cands, survs
