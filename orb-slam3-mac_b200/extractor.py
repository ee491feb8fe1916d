"""Host-side mirror of ORB_SLAM3::ORBextractor (reference include/ORBextractor.h:43-109) over the
C ABI.  Same constructor arguments, same getters, same call semantics; used by tests/ and
bench.py.  The C++ drop-in with the identical class shape is orb-slam3-mac_b200/host/ORBextractor.h.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbxConfig, OrbxError, check  # noqa: F401


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) -- src/ORBextractor.cc:408.

    Extra keyword arguments size the device-resident buffers (the reference allocates per frame):
    max_size=(h, w), max_batch, n_slots, device.
    """

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int,
                 minThFAST: int, *, max_size=(480, 752), max_batch: int = 1, n_slots: int = 2,
                 device: int = 0):
        self.lib = _lib.load()
        cfg = OrbxConfig(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_size[1],
                         max_size[0], max_batch, n_slots, device, 0)
        self.h = C.c_void_p()
        check(self.lib.orbx_create(C.byref(cfg), C.byref(self.h)))
        self.nfeatures, self.scaleFactor, self.nlevels = nfeatures, scaleFactor, nlevels
        self.max_batch, self.n_slots = max_batch, n_slots
        self._keep = {}
        self._channels = 1

    def close(self):
        if getattr(self, "h", None):
            self.lib.orbx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- input stage (N4): cvtColor / cv::remap in front of level 0 ---------------------------------
    FORMATS = {"gray": (0, 1), "bgr": (1, 3), "rgb": (2, 3), "bgra": (3, 4), "rgba": (4, 4)}

    def set_input_format(self, fmt: str = "gray", gray_bits: int = 15):
        """What Tracking::GrabImage* does with a colour image (src/Tracking.cc:354-379): later calls
        take H x W x C images.  gray_bits: 15 = OpenCV >= 4 luma coefficients, 14 = OpenCV 2/3."""
        code, ch = self.FORMATS[fmt]
        check(self.lib.orbx_set_input_format(self.h, code, gray_bits))
        self._channels = ch

    def set_rectify_maps(self, map_x=None, map_y=None):
        """cv::remap(im, imRect, M1, M2, INTER_LINEAR) of Examples/Stereo/stereo_euroc.cc:170-171 in
        front of level 0; maps are the CV_32FC1 outputs of cv::initUndistortRectifyMap."""
        if map_x is None:
            check(self.lib.orbx_set_rectify_maps(self.h, None, None, 0, 0, 0))
            return
        mx = np.ascontiguousarray(map_x, np.float32)
        my = np.ascontiguousarray(map_y, np.float32)
        assert mx.shape == my.shape and mx.ndim == 2
        check(self.lib.orbx_set_rectify_maps(self.h, mx.ctypes.data, my.ctypes.data, mx.shape[1], mx.shape[0],
                                             mx.shape[1]))

    # ---- getters (include/ORBextractor.h:61-81) ------------------------------------------------
    def GetLevels(self) -> int:
        return self.nlevels

    def GetScaleFactor(self) -> float:
        return self.scaleFactor

    def _tables(self):
        n = self.nlevels
        a, b, c, d = (np.zeros(n, np.float32) for _ in range(4))
        e = np.zeros(n, np.int32)
        check(self.lib.orbx_get_tables(self.h, a.ctypes.data, b.ctypes.data, c.ctypes.data,
                                       d.ctypes.data, e.ctypes.data))
        return a, b, c, d, e

    def GetScaleFactors(self):
        return self._tables()[0]

    def GetInverseScaleFactors(self):
        return self._tables()[1]

    def GetScaleSigmaSquares(self):
        return self._tables()[2]

    def GetInverseScaleSigmaSquares(self):
        return self._tables()[3]

    def features_per_level(self):
        return self._tables()[4]

    def max_keypoints(self) -> int:
        return int(self.lib.orbx_max_keypoints(self.h))

    # ---- operator() ------------------------------------------------------------------------------
    def __call__(self, image: np.ndarray | None, vLappingArea=(0, 0), cap: int | None = None):
        """Returns (monoIndex, keypoints[KP_DTYPE], descriptors uint8[n,32]).

        An empty image returns (-1, [], []) like the reference (src/ORBextractor.cc:1072-1073);
        a non-uint8 image raises (the reference asserts CV_8UC1, :1076)."""
        if image is None or image.size == 0:
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        ch = self._channels
        if image.dtype != np.uint8 or (ch == 1 and image.ndim != 2) or (ch > 1 and (image.ndim != 3 or image.shape[2] != ch)):
            raise OrbxError(_lib.ORBX_E_BAD_TYPE, "image must be 8-bit with %d channel(s)" % ch)
        if image.strides[1] != ch or (ch > 1 and image.strides[2] != 1):
            image = np.ascontiguousarray(image)
        cap = cap or self.max_keypoints()
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n, mono = C.c_int(0), C.c_int(0)
        check(self.lib.orbx_extract(self.h, image.ctypes.data, image.shape[1], image.shape[0],
                                    image.strides[0], int(vLappingArea[0]), int(vLappingArea[1]),
                                    kps.ctypes.data, desc.ctypes.data, cap, C.byref(n), C.byref(mono)))
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, lapping=None, cap: int | None = None):
        """images: list of equally-shaped uint8 arrays (or one [B,H,W] array)."""
        imgs = [np.ascontiguousarray(im) for im in images]
        B = len(imgs)
        h, w = imgs[0].shape[:2]
        cap = cap or self.max_keypoints()
        ptrs = (C.c_void_p * B)(*[im.ctypes.data for im in imgs])
        lap = None
        if lapping is not None:
            lap = np.ascontiguousarray(np.asarray(lapping, np.int32).reshape(B, 2))
        kps = np.zeros((B, cap), KP_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        mono = np.zeros(B, np.int32)
        status = np.zeros(B, np.int32)
        check(self.lib.orbx_extract_batch(self.h, ptrs, B, w, h, imgs[0].strides[0],
                                          lap.ctypes.data if lap is not None else None,
                                          kps.ctypes.data, desc.ctypes.data, cap, n.ctypes.data,
                                          mono.ctypes.data, status.ctypes.data))
        return [(int(mono[i]), kps[i, :n[i]].copy(), desc[i, :n[i]].copy()) for i in range(B)]

    # ---- mvImagePyramid (include/ORBextractor.h:83) ------------------------------------------------
    def level_size(self, w: int, h: int, level: int):
        lw, lh = C.c_int(), C.c_int()
        check(self.lib.orbx_level_size(self.h, w, h, level, C.byref(lw), C.byref(lh)))
        return lw.value, lh.value

    def pyramid_level(self, level: int, shape, frame: int = 0, slot: int = 0, pad: int = 0):
        lw, lh = self.level_size(shape[1], shape[0], level)
        out = np.zeros((lh + 2 * pad, lw + 2 * pad), np.uint8)
        check(self.lib.orbx_get_pyramid_level(self.h, slot, frame, level, pad, out.ctypes.data,
                                              out.strides[0]))
        return out

    def device_pyramid(self, frame: int = 0, slot: int = 0):
        """orbx_pyramid_view of the last run: device pointers for a same-GPU consumer
        (ORBmatcher.ComputeStereoMatches)."""
        v = _lib.OrbxPyramidView()
        check(self.lib.orbx_device_pyramid(self.h, slot, frame, C.byref(v)))
        return v

    # ---- stage taps ---------------------------------------------------------------------------------
    def debug_candidates(self, level: int, frame: int = 0, slot: int = 0):
        n = self.lib.orbx_debug_candidates(self.h, slot, frame, level, None, 0)
        if n < 0:
            check(n)
        out = np.zeros((n, 3), np.int32)
        if n:
            self.lib.orbx_debug_candidates(self.h, slot, frame, level, out.ctypes.data, n)
        return out

    def debug_survivors(self, level: int, frame: int = 0, slot: int = 0):
        n = self.lib.orbx_debug_survivors(self.h, slot, frame, level, None, 0)
        if n < 0:
            check(n)
        out = np.zeros((n, 3), np.int32)
        if n:
            self.lib.orbx_debug_survivors(self.h, slot, frame, level, out.ctypes.data, n)
        return out

    def launch_count(self, slot: int = 0) -> int:
        return int(self.lib.orbx_last_launch_count(self.h, slot))
