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
MREF_SO = os.path.join(HERE, "_ref", "liborbmref.so")
MDROP_SO = os.path.join(HERE, "_ref", "liborbmdrop.so")   # the same wrapper over the PRODUCT's C++ ORBmatcher drop-in
PORT_SO = os.path.join(HERE, "_build", "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build(target: str = "all") -> None:
    subprocess.run(["make", "-s", "-f", os.path.join(HERE, "Makefile"), target], check=True)


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def have_mref() -> bool:
    return os.path.exists(MREF_SO)


def have_mdrop() -> bool:
    return os.path.exists(MDROP_SO)


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


# ---- matcher oracle --------------------------------------------------------------------------------
class _Renamed:
    """liborbmref.so exports the orbo_* surface under the prefix orbr_."""

    def __init__(self, lib):
        self._lib = lib

    def __getattr__(self, name):
        return getattr(self._lib, "orbr_" + name[len("orbo_"):] if name.startswith("orbo_") else name)


def canon_inv_z(a):
    """inv_z values that survive float(1 / float(1 / inv_z)): what SearchByProjection(Cur, Last) computes from a
    depth (src/ORBmatcher.cc:2001).  The reference-held oracle refuses other inputs."""
    a = np.asarray(a, np.float32).copy()
    for _ in range(4):
        z = (1.0 / a.astype(np.float64)).astype(np.float32)
        a = (1.0 / z.astype(np.float64)).astype(np.float32)
    return a


class MatchOracle:
    """The reference's Hamming searches on the same POD snapshots the product takes (struct layouts of
    include/orbm.h).  Frame objects are orb_slam3_mac_b200.matcher.Frame.
      kind="port": oracle/match_oracle.cpp, our restatement (liborb_oracle.so);
      kind="ref":  the reference's own src/ORBmatcher.cc / Frame.cc / MapPoint.cc / DBoW2, compiled verbatim
                   (oracle/_ref/liborbmref.so through oracle/ref_match_capi.cpp)."""

    def __init__(self, kind: str = "port"):
        self.kind = kind
        if kind == "ref":
            if not have_mref():
                build("mref")
            self.lib = _Renamed(C.CDLL(MREF_SO))
        elif kind == "drop":
            # NOT an oracle: the reference's Frame / MapPoint objects driven through the product's C++ class
            # orb-slam3-mac_b200/host/ORBmatcher.{h,cc} (-> liborbx_b200.so, needs a GPU); tests/test_cpp_dropin.py
            # compares it with kind="ref".  Tracking searches only.
            if not have_mdrop():
                build("mdrop")
            self.lib = _Renamed(C.CDLL(MDROP_SO))
        else:
            if not have_port():
                build("port")
            self.lib = C.CDLL(PORT_SO)

    def distance(self, a, b):
        return self.lib.orbo_descriptor_distance(C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data))

    def features_in_area(self, F, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.zeros(max(1, F.N), np.int32)
        v = F.view()
        n = self.lib.orbo_features_in_area(C.byref(v), C.c_float(x), C.c_float(y), C.c_float(r), minLevel, maxLevel,
                                           C.c_void_p(out.ctypes.data), len(out))
        return out[:n].copy()

    def search_by_projection_points(self, F, mps, mp_desc, th, far, th_far, nnratio, kp_obs=None):
        obs = np.full(F.N, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(F.N, np.int32)
        v = F.view()
        n = self.lib.orbo_search_by_projection_points(
            C.byref(v), C.c_void_p(mps.ctypes.data), C.c_void_p(mp_desc.ctypes.data), len(mps), C.c_float(th),
            int(far), C.c_float(th_far), C.c_float(nnratio), C.c_void_p(obs.ctypes.data), C.c_void_p(match.ctypes.data))
        return n, match, obs

    def search_by_projection_frames(self, cur, last_obs, last_desc, th, fwd, bwd, mbf, check_ori, kp_obs=None):
        obs = np.full(cur.N, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(cur.N, np.int32)
        v = cur.view()
        n = self.lib.orbo_search_by_projection_frames(
            C.byref(v), C.c_void_p(last_obs.ctypes.data), C.c_void_p(last_desc.ctypes.data), len(last_obs),
            C.c_float(th), int(fwd), int(bwd), C.c_float(mbf), int(check_ori), C.c_void_p(obs.ctypes.data),
            C.c_void_p(match.ctypes.data))
        return n, match, obs

    def search_by_projection_points_stereo(self, F, R, mps, mps_r, mp_desc, th, far, th_far, nnratio, kp_obs=None):
        n_all = F.N + R.N
        obs = np.full(n_all, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(n_all, np.int32)
        v, r = F.view(), R.view()
        n = self.lib.orbo_search_by_projection_points_stereo(
            C.byref(v), C.byref(r), C.c_void_p(mps.ctypes.data), C.c_void_p(mps_r.ctypes.data),
            C.c_void_p(mp_desc.ctypes.data), len(mps), C.c_float(th), int(far), C.c_float(th_far), C.c_float(nnratio),
            C.c_void_p(obs.ctypes.data), C.c_void_p(match.ctypes.data))
        return n, match, obs

    def search_by_projection_frames_stereo(self, cur, R, last_obs, uv_right, last_desc, th, fwd, bwd, check_ori,
                                           kp_obs=None):
        n_all = cur.N + R.N
        obs = np.full(n_all, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(n_all, np.int32)
        uvr = np.ascontiguousarray(uv_right, np.float32)
        v, r = cur.view(), R.view()
        n = self.lib.orbo_search_by_projection_frames_stereo(
            C.byref(v), C.byref(r), C.c_void_p(last_obs.ctypes.data), C.c_void_p(uvr.ctypes.data),
            C.c_void_p(last_desc.ctypes.data), len(last_obs), C.c_float(th), int(fwd), int(bwd), int(check_ori),
            C.c_void_p(obs.ctypes.data), C.c_void_p(match.ctypes.data))
        return n, match, obs

    def search_by_bow(self, KF, kf_mp_ok, F, nleft, nnratio, check_ori):
        ok = np.ascontiguousarray(kf_mp_ok, np.uint8)
        matches = np.zeros(len(F.keys), np.int32)
        kv, fv = KF.view(), F.view()
        n = self.lib.orbo_search_by_bow(C.byref(kv), C.c_void_p(ok.ctypes.data), C.byref(fv), int(nleft),
                                        C.c_float(nnratio), int(check_ori), C.c_void_p(matches.ctypes.data))
        return n, matches

    def search_for_initialization(self, F1, F2, prev, window, nnratio, check_ori):
        prev = np.ascontiguousarray(prev, np.float32).reshape(-1, 2).copy()
        m12 = np.zeros(F1.N, np.int32)
        v1, v2 = F1.view(), F2.view()
        n = self.lib.orbo_search_for_initialization(C.byref(v1), C.byref(v2), C.c_void_p(prev.ctypes.data),
                                                    C.c_void_p(m12.ctypes.data), int(window), C.c_float(nnratio),
                                                    int(check_ori))
        return n, m12, prev

    def stereo_rows(self, left, right, n_rows, min_d, max_d):
        if self.kind == "ref":
            raise NotImplementedError("the reference has no separate row search: use compute_stereo_matches")
        bi = np.zeros(left.N, np.int32)
        bd = np.zeros(left.N, np.int32)
        self.lib.orbo_stereo_rows(C.c_void_p(left.keys.ctypes.data), C.c_void_p(left.desc.ctypes.data), left.N,
                                  C.c_void_p(right.keys.ctypes.data), C.c_void_p(right.desc.ctypes.data), right.N,
                                  C.c_void_p(left.scale_factors.ctypes.data), int(n_rows), C.c_float(min_d),
                                  C.c_float(max_d), C.c_void_p(bi.ctypes.data), C.c_void_p(bd.ctypes.data))
        return bi, bd

    def compute_stereo_matches(self, left, right, pyr_l, pyr_r, mb, mbf):
        """pyr_l / pyr_r: per level a uint8 array padded by 19 px (OracleExtractor.level(l, 19))."""
        nl = len(pyr_l)
        keep = [np.ascontiguousarray(a) for a in list(pyr_l) + list(pyr_r)]
        pl = (C.c_void_p * nl)(*[a.ctypes.data + 19 * a.strides[0] + 19 for a in keep[:nl]])
        pr = (C.c_void_p * nl)(*[a.ctypes.data + 19 * a.strides[0] + 19 for a in keep[nl:]])
        sl = (C.c_int * nl)(*[a.strides[0] for a in keep[:nl]])
        sr = (C.c_int * nl)(*[a.strides[0] for a in keep[nl:]])
        cr = (C.c_int * nl)(*[a.shape[1] - 38 for a in keep[nl:]])
        inv = np.ascontiguousarray(1.0 / left.scale_factors.astype(np.float32), np.float32)
        u = np.zeros(left.N, np.float32)
        d = np.zeros(left.N, np.float32)
        n = self.lib.orbo_compute_stereo_matches(
            C.c_void_p(left.keys.ctypes.data), C.c_void_p(left.desc.ctypes.data), left.N,
            C.c_void_p(right.keys.ctypes.data), C.c_void_p(right.desc.ctypes.data), right.N,
            C.c_void_p(left.scale_factors.ctypes.data), C.c_void_p(inv.ctypes.data), pl, sl, pr, sr, cr,
            int(keep[0].shape[0] - 38), C.c_float(mb), C.c_float(mbf), C.c_void_p(u.ctypes.data),
            C.c_void_p(d.ctypes.data), *([nl] if self.kind == "ref" else []))
        return u, d, n

    # ---- SURVEY 8(f) N3 ----------------------------------------------------------------------------
    @staticmethod
    def _p(a):
        return None if a is None else C.c_void_p(a.ctypes.data)

    def search_by_bow_kf(self, KF1, ok1, KF2, ok2, n_un1, n_un2, nnratio, check_ori):
        ok1, ok2 = np.ascontiguousarray(ok1, np.uint8), np.ascontiguousarray(ok2, np.uint8)
        m12 = np.zeros(len(KF1.keys), np.int32)
        v1, v2 = KF1.view(), KF2.view()
        n = self.lib.orbo_search_by_bow_kf(C.byref(v1), self._p(ok1), int(n_un1), C.byref(v2), self._p(ok2), int(n_un2),
                                           C.c_float(nnratio), int(check_ori), self._p(m12))
        return n, m12

    def search_by_projection_sim3(self, KF, pts, mp_desc, th, ratio, kp_taken=None):
        taken = np.zeros(KF.N, np.uint8) if kp_taken is None else np.ascontiguousarray(kp_taken, np.uint8)
        match = np.zeros(KF.N, np.int32)
        v = KF.view()
        n = self.lib.orbo_search_by_projection_sim3(C.byref(v), self._p(pts), self._p(mp_desc), len(pts), int(th),
                                                    C.c_float(ratio), self._p(taken), self._p(match))
        return n, match

    def search_by_projection_reloc(self, F, pts, mp_desc, th, orb_dist, check_ori, kp_taken=None):
        taken = np.zeros(F.N, np.uint8) if kp_taken is None else np.ascontiguousarray(kp_taken, np.uint8)
        match = np.zeros(F.N, np.int32)
        v = F.view()
        n = self.lib.orbo_search_by_projection_reloc(C.byref(v), self._p(pts), self._p(mp_desc), len(pts), C.c_float(th),
                                                     int(orb_dist), int(check_ori), self._p(taken), self._p(match))
        return n, match

    def fuse(self, KF, pts, mp_desc, th, inv_level_sigma2=None, idx_offset=0, depth=None, bf=0.0):
        """depth / bf: the camera-frame depth of every point and the keyframe's mbf, from which the caller derived
        pts["ur"] = u - bf * (1 / depth) (src/ORBmatcher.cc:1490); only the reference-held oracle needs them."""
        inv = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        bi = np.zeros(len(pts), np.int32)
        bd = np.zeros(len(pts), np.int32)
        v = KF.view()
        extra = []
        if self.kind == "ref":
            dz = None if depth is None else np.ascontiguousarray(depth, np.float32)
            assert inv is None or dz is not None, "the reference-held Fuse needs the points' depths"
            extra = [self._p(dz), C.c_float(bf)]
        n = self.lib.orbo_fuse(C.byref(v), self._p(pts), self._p(mp_desc), len(pts), C.c_float(th), self._p(inv),
                               int(idx_offset), self._p(bi), self._p(bd), *extra)
        assert n >= 0, n
        return n, bi, bd

    def search_by_sim3(self, KF1, KF2, p1, d1, p2, d2, th):
        m12 = np.zeros(KF1.N, np.int32)
        v1, v2 = KF1.view(), KF2.view()
        n = self.lib.orbo_search_by_sim3(C.byref(v1), C.byref(v2), self._p(p1), self._p(d1), self._p(p2), self._p(d2),
                                         C.c_float(th), self._p(m12))
        return n, m12

    def search_for_triangulation(self, KF1, has1, st1, KF2, has2, st2, only_stereo, ep, sf2, pair_ok, F12, ls2,
                                 check_ori, overload=1):
        """overload 1: SearchForTriangulation(.., bOnlyStereo, bCoarse) src/ORBmatcher.cc:969; 2: the matchAndtriangulate
        form :1212 (no stereo / epipole gates there: pass st1 = st2 = ep = None)."""
        m12 = np.zeros(len(KF1.keys), np.int32)
        v1, v2 = KF1.view(), KF2.view()
        n = self.lib.orbo_search_for_triangulation(C.byref(v1), self._p(has1), self._p(st1), C.byref(v2), self._p(has2),
                                                   self._p(st2), int(only_stereo), self._p(ep), self._p(sf2),
                                                   self._p(pair_ok), self._p(F12), self._p(ls2), int(check_ori),
                                                   self._p(m12), *([int(overload)] if self.kind == "ref" else []))
        assert n >= 0, n
        return n, m12

    def distinctive_descriptors(self, desc, group_ptr):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        gp = np.ascontiguousarray(group_ptr, np.int32)
        best = np.zeros(len(gp) - 1, np.int32)
        self.lib.orbo_distinctive_descriptors(self._p(d), self._p(gp), len(gp) - 1, self._p(best))
        return best

    def bow_transform(self, voc, desc, levelsup):
        """voc: orb_slam3_mac_b200.matcher.Vocabulary (host arrays only).  Returns (bow_ids, bow_vals,
        fv_node_ids, fv_ptr, fv_items)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d)
        bi, bv = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        fn, fp, fi = np.zeros(max(n, 1), np.uint32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.uint32)
        nw, nf = C.c_int32(), C.c_int32()
        self.lib.orbo_bow_transform(len(voc.weight), self._p(voc.child_ptr), self._p(voc.child_ids), self._p(voc.node_desc),
                                    self._p(voc.weight), self._p(voc.word_id), int(voc.L), int(voc.weighting),
                                    int(voc.norm), self._p(d), n, int(levelsup), self._p(bi), self._p(bv), C.byref(nw),
                                    self._p(fn), self._p(fp), self._p(fi), C.byref(nf))
        return bi[:nw.value].copy(), bv[:nw.value].copy(), fn[:nf.value].copy(), fp[:nf.value + 1].copy(), \
            fi[:fp[nf.value]].copy()

    def knn2(self, q, t):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
        idx = np.zeros((len(q), 2), np.int32)
        dist = np.zeros((len(q), 2), np.int32)
        if self.kind == "ref":   # the reference calls cv::BFMatcher (src/Frame.cc:1146): here the shim's restatement of it
            fd = np.zeros((len(q), 2), np.float32)
            self.lib.orbo_shim_knn2(C.c_void_p(q.ctypes.data), len(q), C.c_void_p(t.ctypes.data), len(t),
                                    C.c_void_p(idx.ctypes.data), C.c_void_p(fd.ctypes.data))
            return idx, fd.astype(np.int32)
        self.lib.orbo_knn2(C.c_void_p(q.ctypes.data), len(q), C.c_void_p(t.ctypes.data), len(t),
                           C.c_void_p(idx.ctypes.data), C.c_void_p(dist.ctypes.data))
        return idx, dist


# ---- N4 input-stage primitives (oracle/cvprims.h: cvt_gray_u8, remap_linear_u8) ------------------------
def cvt_gray(img: np.ndarray, rgb_order: bool, bits: int = 15) -> np.ndarray:
    """cvtColor(img, CV_{BGR,RGB,BGRA,RGBA}2GRAY), src/Tracking.cc:354-379."""
    lib = C.CDLL(PORT_SO)
    img = np.ascontiguousarray(img, np.uint8)
    h, w, ch = img.shape
    out = np.zeros((h, w), np.uint8)
    lib.orbo_prim_cvt_gray(C.c_void_p(img.ctypes.data), h, w, C.c_size_t(img.strides[0]), ch, int(rgb_order), bits,
                           C.c_void_p(out.ctypes.data), C.c_size_t(out.strides[0]))
    return out


def remap_linear(img: np.ndarray, map_x: np.ndarray, map_y: np.ndarray) -> np.ndarray:
    """cv::remap(img, out, map_x, map_y, INTER_LINEAR), Examples/Stereo/stereo_euroc.cc:170-171."""
    lib = C.CDLL(PORT_SO)
    img = np.ascontiguousarray(img, np.uint8)
    mx, my = np.ascontiguousarray(map_x, np.float32), np.ascontiguousarray(map_y, np.float32)
    ch = 1 if img.ndim == 2 else img.shape[2]
    h, w = mx.shape
    out = np.zeros((h, w) if ch == 1 else (h, w, ch), np.uint8)
    lib.orbo_prim_remap(C.c_void_p(img.ctypes.data), img.shape[0], img.shape[1], C.c_size_t(img.strides[0]), ch,
                        C.c_void_p(mx.ctypes.data), C.c_void_p(my.ctypes.data), C.c_void_p(out.ctypes.data), h, w,
                        C.c_size_t(out.strides[0]))
    return out
