"""Host-side mirror of the reference's Hamming searches over the C ABI of include/orbm.h.

`Frame` is the POD snapshot of the parts of ORB_SLAM3::Frame the searches read (keypoints,
descriptors, image bounds, grid pitch, scale factors; include/Frame.h); `ORBmatcher` carries
mfNNratio / mbCheckOrientation like the reference class (include/ORBmatcher.h:36-44) and exposes
the searches with the reference's names.  All distances run on the GPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, check
from ._lib_match import (AREA_QUERY_DTYPE, BowView, CAMERA_DTYPE, FrameRight, FrameView, LAST_OBS_DTYPE, MAP_POINT_DTYPE,  # noqa: F401
                         MAP_POINT_RIGHT_DTYPE, PAIR_TEST, PROJ_POINT_DTYPE)

FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48


class Frame:
    """Snapshot of a Frame: mvKeysUn, mDescriptors, mnMinX.., mfGridElement*Inv, mvScaleFactors, mvuRight."""

    def __init__(self, keys, desc, width, height, scale_factors, u_right=None, bounds=None):
        self.keys = np.ascontiguousarray(keys, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        assert len(self.keys) == len(self.desc)
        self.scale_factors = np.ascontiguousarray(scale_factors, np.float32)
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        # no distortion: bounds are the image (src/Frame.cc:773-800, ComputeImageBounds)
        mnx, mny, mxx, mxy = bounds if bounds is not None else (0.0, 0.0, float(width), float(height))
        self.bounds = (np.float32(mnx), np.float32(mny), np.float32(mxx), np.float32(mxy))
        # src/Frame.cc:119-120: static_cast<float>(FRAME_GRID_COLS) / (mnMaxX - mnMinX)
        self.grid_w_inv = np.float32(FRAME_GRID_COLS) / (self.bounds[2] - self.bounds[0])
        self.grid_h_inv = np.float32(FRAME_GRID_ROWS) / (self.bounds[3] - self.bounds[1])

    @property
    def N(self):
        return len(self.keys)

    def view(self) -> FrameView:
        return FrameView(self.keys.ctypes.data, self.desc.ctypes.data, len(self.keys), self.bounds[0],
                         self.bounds[1], self.bounds[2], self.bounds[3], self.grid_w_inv, self.grid_h_inv,
                         self.scale_factors.ctypes.data, len(self.scale_factors),
                         None if self.u_right is None else self.u_right.ctypes.data)


class ResidentFrame:
    """orbm_frame: a frame whose keypoints, descriptors and grid stay on the matcher's device across searches.
    Takes the place of a Frame in every ORBmatcher call (view() returns the library's own view, whose address is
    how the library recognises the frame)."""

    def __init__(self, matcher: "ORBmatcher", frame: "Frame" = None, *, d_keys: int = 0, d_desc: int = 0, n: int = 0,
                 camera=None, width: int = 0, height: int = 0, scale_factors=None):
        self.lib, self._m = matcher.lib, matcher
        self.h = C.c_void_p()
        if frame is not None:
            v = frame.view()
            check(self.lib.orbm_frame_upload(matcher.ctx, C.byref(v), C.byref(self.h)))
        else:
            sf = np.ascontiguousarray(scale_factors, np.float32)
            cam = None
            if camera is not None:
                cam = np.zeros(1, CAMERA_DTYPE)
                for k, val in zip(CAMERA_DTYPE.names, camera):
                    cam[k] = val
            check(self.lib.orbm_frame_from_device(matcher.ctx, C.c_void_p(d_keys), C.c_void_p(d_desc), int(n),
                                                  None if cam is None else cam.ctypes.data, int(width), int(height),
                                                  sf.ctypes.data, len(sf), C.byref(self.h)))
        self._vp = self.lib.orbm_frame_get_view(self.h)
        v = self.view()
        n = v.n
        self.keys = np.ctypeslib.as_array(C.cast(v.keys, C.POINTER(C.c_uint8)), shape=(max(n, 1) * 28,))[:n * 28].view(KP_DTYPE) \
            if n else np.zeros(0, KP_DTYPE)
        self.desc = np.ctypeslib.as_array(C.cast(v.desc, C.POINTER(C.c_uint8)), shape=(max(n, 1) * 32,))[:n * 32].reshape(-1, 32) \
            if n else np.zeros((0, 32), np.uint8)
        self.bounds = (np.float32(v.min_x), np.float32(v.min_y), np.float32(v.max_x), np.float32(v.max_y))
        self.grid_w_inv, self.grid_h_inv = np.float32(v.grid_w_inv), np.float32(v.grid_h_inv)
        self.scale_factors = np.ctypeslib.as_array(C.cast(v.scale_factors, C.POINTER(C.c_float)), shape=(v.nlevels,))
        self.u_right = None

    @property
    def N(self):
        return len(self.keys)

    def view(self) -> FrameView:
        return FrameView.from_address(self._vp)

    def set_u_right(self, u_right):
        self.u_right = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        check(self.lib.orbm_frame_set_u_right(self.h, None if self.u_right is None else self.u_right.ctypes.data))

    def bank(self) -> "DescriptorBank":
        """This frame's descriptors as a bank (row = keypoint index)."""
        return DescriptorBank(self._m, _view_of=self)

    def close(self):
        if getattr(self, "h", None) and self._m.ctx:
            self.lib.orbm_frame_destroy(self.h)
        self.h = None


class DescriptorBank:
    """orbm_bank: 32-byte descriptor rows kept on the matcher's device (the map points' mDescriptor, one row per
    MapPoint; or a resident frame's descriptors).  Pass it where a search takes descriptors, with `rows`."""

    def __init__(self, matcher: "ORBmatcher", capacity: int = 0, *, _view_of: "ResidentFrame" = None):
        self.lib, self._m, self._owner = matcher.lib, matcher, _view_of
        if _view_of is not None:
            self.h = C.c_void_p(self.lib.orbm_frame_bank(_view_of.h))
            self._owned = False
        else:
            self.h = C.c_void_p()
            check(self.lib.orbm_bank_create(matcher.ctx, int(capacity), C.byref(self.h)))
            self._owned = True

    @property
    def rows(self) -> int:
        return int(self.lib.orbm_bank_rows(self.h))

    def write(self, first_row: int, desc):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        check(self.lib.orbm_bank_write(self.h, int(first_row), len(d), d.ctypes.data))

    def close(self):
        if getattr(self, "_owned", False) and self.h and self._m.ctx:
            self.lib.orbm_bank_destroy(self.h)
        self.h = None


class BowSide:
    """One side of SearchByBoW: keypoints, descriptors and the DBoW2::FeatureVector as
    {node id: [feature indices]} (fisheye rigs: left ++ right, as mDescriptors is laid out)."""

    def __init__(self, keys, desc, feat_vec: dict):
        self.keys = np.ascontiguousarray(keys, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        ids = sorted(feat_vec)
        self.node_ids = np.array(ids, np.uint32)
        self.node_ptr = np.zeros(len(ids) + 1, np.int32)
        items = []
        for k, nid in enumerate(ids):
            items.extend(int(v) for v in feat_vec[nid])
            self.node_ptr[k + 1] = len(items)
        self.node_items = np.array(items, np.uint32)

    def view(self) -> BowView:
        return BowView(self.keys.ctypes.data, self.desc.ctypes.data, len(self.keys), len(self.node_ids),
                       self.node_ids.ctypes.data, self.node_ptr.ctypes.data, self.node_items.ctypes.data)


class Vocabulary:
    """DBoW2::TemplatedVocabulary<FORB> as POD (include/orbm.h, orbm_vocab_create): children as CSR in the order of
    Node::children, node descriptors, word weights and ids; weighting 0 = TF_IDF, norm 1 = L1 (ORBvoc)."""

    def __init__(self, child_ptr, child_ids, node_desc, weight, word_id, L, weighting=0, norm=1):
        self.child_ptr = np.ascontiguousarray(child_ptr, np.int32)
        self.child_ids = np.ascontiguousarray(child_ids, np.uint32)
        self.node_desc = np.ascontiguousarray(node_desc, np.uint8).reshape(-1, 32)
        self.weight = np.ascontiguousarray(weight, np.float64)
        self.word_id = np.ascontiguousarray(word_id, np.uint32)
        self.L, self.weighting, self.norm = int(L), int(weighting), int(norm)
        assert len(self.child_ptr) == len(self.weight) + 1 == len(self.node_desc) + 1


class RightHalf:
    """The right camera's half of a fisheye stereo Frame (Nleft != -1): mvKeysRight, the descriptor rows
    Nleft.., mvLeftToRightMatch, mvRightToLeftMatch."""

    def __init__(self, keys, desc, left_to_right=None, right_to_left=None):
        self.keys = np.ascontiguousarray(keys, KP_DTYPE)
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        self.l2r = None if left_to_right is None else np.ascontiguousarray(left_to_right, np.int32)
        self.r2l = None if right_to_left is None else np.ascontiguousarray(right_to_left, np.int32)

    @property
    def N(self):
        return len(self.keys)

    def view(self) -> FrameRight:
        return FrameRight(self.keys.ctypes.data, self.desc.ctypes.data, len(self.keys),
                          None if self.l2r is None else self.l2r.ctypes.data,
                          None if self.r2l is None else self.r2l.ctypes.data)


class ORBmatcher:
    TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30

    def __init__(self, nnratio: float = 0.6, checkOri: bool = True, device: int = 0):
        self.lib = _lib.load()
        self.mfNNratio, self.mbCheckOrientation = float(nnratio), bool(checkOri)
        self.ctx = C.c_void_p()
        check(self.lib.orbm_create(device, C.byref(self.ctx)))

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.orbm_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def area_candidates(self, F: "Frame", queries, q_desc=None, cap=None):
        """Device-side AssignFeaturesToGrid + GetFeaturesInArea for a batch of window queries
        (src/Frame.cc:377-408, 645-726): returns (row_ptr, cand, dist); dist is None without q_desc."""
        q = np.ascontiguousarray(queries, AREA_QUERY_DTYPE)
        nq = len(q)
        row_ptr = np.zeros(nq + 1, np.int32)
        qd = None if q_desc is None else np.ascontiguousarray(q_desc, np.uint8).reshape(-1, 32)
        cap = int(cap) if cap is not None else max(1, nq * max(1, F.N))
        cand = np.zeros(cap, np.int32)
        dist = np.zeros(cap, np.uint16)
        v = F.view()
        n = self.lib.orbm_area_candidates(self.ctx, C.byref(v), q.ctypes.data, nq,
                                          None if qd is None else qd.ctypes.data, 0 if qd is None else len(qd),
                                          row_ptr.ctypes.data, cand.ctypes.data, dist.ctypes.data, cap)
        if n < 0:
            check(n)
        return row_ptr, cand[:n].copy(), (None if qd is None else dist[:n].copy())

    # ---- layer 1 ------------------------------------------------------------------------------------
    def hamming_csr(self, q, t, row_ptr, cand):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
        row_ptr = np.ascontiguousarray(row_ptr, np.int32)
        cand = np.ascontiguousarray(cand, np.int32)
        out = np.zeros(len(cand), np.uint16)
        check(self.lib.orbm_hamming_csr(self.ctx, q.ctypes.data, len(q), t.ctypes.data, len(t),
                                        row_ptr.ctypes.data, cand.ctypes.data, len(cand), out.ctypes.data))
        return out

    def knn2(self, q, t):
        """BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2) -> (idx[nq,2], dist[nq,2])."""
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
        idx = np.zeros((len(q), 2), np.int32)
        dist = np.zeros((len(q), 2), np.int32)
        check(self.lib.orbm_knn2(self.ctx, q.ctypes.data, len(q), t.ctypes.data, len(t), idx.ctypes.data,
                                 dist.ctypes.data))
        return idx, dist

    def stereo_rows(self, left: Frame, right: Frame, min_d: float, max_d: float):
        """Hamming part of Frame::ComputeStereoMatches -> (best right index or -1, best distance)."""
        bi = np.zeros(left.N, np.int32)
        bd = np.zeros(left.N, np.int32)
        check(self.lib.orbm_stereo_rows(self.ctx, left.keys.ctypes.data, left.desc.ctypes.data, left.N,
                                        right.keys.ctypes.data, right.desc.ctypes.data, right.N,
                                        left.scale_factors.ctypes.data, len(left.scale_factors), min_d, max_d,
                                        bi.ctypes.data, bd.ctypes.data))
        return bi, bd

    def ComputeStereoMatches(self, left: Frame, right: Frame, pyr_left, pyr_right, mb: float, mbf: float):
        """Frame::ComputeStereoMatches (src/Frame.cc:802-980) -> (mvuRight, mvDepth).  pyr_left /
        pyr_right: ORBextractor.device_pyramid() of the two extractions (device resident)."""
        import ctypes as C
        u = np.zeros(left.N, np.float32)
        d = np.zeros(left.N, np.float32)
        inv = np.ascontiguousarray(1.0 / left.scale_factors.astype(np.float32), np.float32) \
            if getattr(left, "inv_scale_factors", None) is None else left.inv_scale_factors
        check(self.lib.orbm_compute_stereo_matches(
            self.ctx, left.keys.ctypes.data, left.desc.ctypes.data, left.N, right.keys.ctypes.data,
            right.desc.ctypes.data, right.N, left.scale_factors.ctypes.data, inv.ctypes.data,
            len(left.scale_factors), C.byref(pyr_left), C.byref(pyr_right), mb, mbf, u.ctypes.data, d.ctypes.data))
        return u, d

    # ---- layer 2 ------------------------------------------------------------------------------------
    def SearchByProjection(self, F: Frame, map_points, mp_desc, th=1.0, bFarPoints=False, thFarPoints=50.0,
                           kp_obs=None, rows=None):
        """SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints).
        mp_desc: the map points' descriptors [M, 32], or a DescriptorBank holding them on the device (rows = the bank
        row of each map point, None for row i).  Returns (nmatches, kp_match[N], kp_obs[N])."""
        mps = np.ascontiguousarray(map_points, MAP_POINT_DTYPE)
        obs = np.full(F.N, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(F.N, np.int32)
        v = F.view()
        if isinstance(mp_desc, DescriptorBank):
            r = None if rows is None else np.ascontiguousarray(rows, np.int32)
            n = self.lib.orbm_search_by_projection_points_bank(
                self.ctx, C.byref(v), mps.ctypes.data, mp_desc.h, None if r is None else r.ctypes.data, len(mps), th,
                int(bFarPoints), thFarPoints, self.mfNNratio, obs.ctypes.data, match.ctypes.data)
        else:
            mp_desc = np.ascontiguousarray(mp_desc, np.uint8).reshape(-1, 32)
            n = self.lib.orbm_search_by_projection_points(self.ctx, C.byref(v), mps.ctypes.data, mp_desc.ctypes.data,
                                                          len(mps), th, int(bFarPoints), thFarPoints, self.mfNNratio,
                                                          obs.ctypes.data, match.ctypes.data)
        if n < 0:
            check(n)
        return n, match, obs

    def SearchByProjectionFrames(self, CurrentFrame: Frame, last_obs, last_desc, th, bForward=False,
                                 bBackward=False, mbf=0.0, kp_obs=None, rows=None):
        """SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono).
        last_desc: one descriptor per observation [n_last, 32] (its map point's), or a DescriptorBank holding the map
        points' descriptors with rows = the bank row of each observation's map point."""
        lo = np.ascontiguousarray(last_obs, LAST_OBS_DTYPE)
        obs = np.full(CurrentFrame.N, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(CurrentFrame.N, np.int32)
        v = CurrentFrame.view()
        if isinstance(last_desc, DescriptorBank):
            r = None if rows is None else np.ascontiguousarray(rows, np.int32)
            n = self.lib.orbm_search_by_projection_frames_bank(
                self.ctx, C.byref(v), lo.ctypes.data, last_desc.h, None if r is None else r.ctypes.data, len(lo), th,
                int(bForward), int(bBackward), mbf, int(self.mbCheckOrientation), obs.ctypes.data, match.ctypes.data)
        else:
            last_desc = np.ascontiguousarray(last_desc, np.uint8).reshape(-1, 32)
            n = self.lib.orbm_search_by_projection_frames(self.ctx, C.byref(v), lo.ctypes.data, last_desc.ctypes.data,
                                                          len(lo), th, int(bForward), int(bBackward), mbf,
                                                          int(self.mbCheckOrientation), obs.ctypes.data, match.ctypes.data)
        if n < 0:
            check(n)
        return n, match, obs

    def SearchByProjectionStereo(self, F: Frame, R: RightHalf, map_points, map_points_right, mp_desc, th=1.0,
                                 bFarPoints=False, thFarPoints=50.0, kp_obs=None):
        """SearchByProjection(F, vpMapPoints, ...) on a fisheye stereo frame (src/ORBmatcher.cc:48-218 with
        Nleft != -1).  Returns (nmatches, kp_match[Nleft+Nright], kp_obs[Nleft+Nright])."""
        mps = np.ascontiguousarray(map_points, MAP_POINT_DTYPE)
        mpr = np.ascontiguousarray(map_points_right, MAP_POINT_RIGHT_DTYPE)
        mp_desc = np.ascontiguousarray(mp_desc, np.uint8).reshape(-1, 32)
        n_all = F.N + R.N
        obs = np.full(n_all, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(n_all, np.int32)
        v, r = F.view(), R.view()
        n = self.lib.orbm_search_by_projection_points_stereo(
            self.ctx, C.byref(v), C.byref(r), mps.ctypes.data, mpr.ctypes.data, mp_desc.ctypes.data, len(mps), th,
            int(bFarPoints), thFarPoints, self.mfNNratio, obs.ctypes.data, match.ctypes.data)
        if n < 0:
            check(n)
        return n, match, obs

    def SearchByProjectionFramesStereo(self, CurrentFrame: Frame, R: RightHalf, last_obs, uv_right, last_desc, th,
                                       bForward=False, bBackward=False, kp_obs=None):
        """SearchByProjection(CurrentFrame, LastFrame, th, bMono) with a fisheye stereo current frame."""
        lo = np.ascontiguousarray(last_obs, LAST_OBS_DTYPE)
        uvr = np.ascontiguousarray(uv_right, np.float32).reshape(-1, 2)
        last_desc = np.ascontiguousarray(last_desc, np.uint8).reshape(-1, 32)
        n_all = CurrentFrame.N + R.N
        obs = np.full(n_all, -1, np.int32) if kp_obs is None else np.ascontiguousarray(kp_obs, np.int32).copy()
        match = np.zeros(n_all, np.int32)
        v, r = CurrentFrame.view(), R.view()
        n = self.lib.orbm_search_by_projection_frames_stereo(
            self.ctx, C.byref(v), C.byref(r), lo.ctypes.data, uvr.ctypes.data, last_desc.ctypes.data, len(lo), th,
            int(bForward), int(bBackward), int(self.mbCheckOrientation), obs.ctypes.data, match.ctypes.data)
        if n < 0:
            check(n)
        return n, match, obs

    def SearchByBoW(self, KF: BowSide, kf_mp_ok, F: BowSide, Nleft: int = -1):
        """SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches), src/ORBmatcher.cc:273-475.
        Returns (nmatches, matches[F.N] = keyframe feature index or -1)."""
        ok = np.ascontiguousarray(kf_mp_ok, np.uint8)
        matches = np.zeros(len(F.keys), np.int32)
        kv, fv = KF.view(), F.view()
        n = self.lib.orbm_search_by_bow(self.ctx, C.byref(kv), ok.ctypes.data, C.byref(fv), int(Nleft), self.mfNNratio,
                                        int(self.mbCheckOrientation), matches.ctypes.data)
        if n < 0:
            check(n)
        return n, matches

    def SearchForInitialization(self, F1: Frame, F2: Frame, vbPrevMatched, windowSize=10):
        """Returns (nmatches, vnMatches12[N1], updated vbPrevMatched[N1,2])."""
        prev = np.ascontiguousarray(vbPrevMatched, np.float32).reshape(-1, 2).copy()
        m12 = np.zeros(F1.N, np.int32)
        v1, v2 = F1.view(), F2.view()
        n = self.lib.orbm_search_for_initialization(self.ctx, C.byref(v1), C.byref(v2), prev.ctypes.data,
                                                    m12.ctypes.data, int(windowSize), self.mfNNratio,
                                                    int(self.mbCheckOrientation))
        if n < 0:
            check(n)
        return n, m12, prev

    # ---- LocalMapping / LoopClosing / relocalisation searches -------------------------
    @staticmethod
    def _u8(a, n):
        return np.zeros(n, np.uint8) if a is None else np.ascontiguousarray(a, np.uint8)

    def SearchByBoWKF(self, KF1: BowSide, mp_ok1, KF2: BowSide, mp_ok2, n_un1: int = -1, n_un2: int = -1):
        """SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12), src/ORBmatcher.cc:827-967.
        Returns (nmatches, matches12[KF1.N] = KF2 feature index or -1)."""
        ok1, ok2 = self._u8(mp_ok1, len(KF1.keys)), self._u8(mp_ok2, len(KF2.keys))
        m12 = np.zeros(len(KF1.keys), np.int32)
        v1, v2 = KF1.view(), KF2.view()
        n = self.lib.orbm_search_by_bow_kf(self.ctx, C.byref(v1), ok1.ctypes.data, int(n_un1), C.byref(v2),
                                           ok2.ctypes.data, int(n_un2), self.mfNNratio, int(self.mbCheckOrientation),
                                           m12.ctypes.data)
        if n < 0:
            check(n)
        return n, m12

    def SearchByProjectionSim3(self, KF: Frame, points, mp_desc, th: int, ratioHamming: float = 1.0, kp_taken=None):
        """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:477-708
        (both overloads).  points: PROJ_POINT_DTYPE.  Returns (nmatches, kp_match[KF.N])."""
        pts = np.ascontiguousarray(points, PROJ_POINT_DTYPE)
        d = np.ascontiguousarray(mp_desc, np.uint8).reshape(-1, 32)
        taken = self._u8(kp_taken, KF.N)
        match = np.zeros(KF.N, np.int32)
        v = KF.view()
        n = self.lib.orbm_search_by_projection_sim3(self.ctx, C.byref(v), pts.ctypes.data, d.ctypes.data, len(pts),
                                                    int(th), float(ratioHamming), taken.ctypes.data, match.ctypes.data)
        if n < 0:
            check(n)
        return n, match

    def SearchByProjectionReloc(self, F: Frame, points, mp_desc, th: float, ORBdist: int, kp_taken=None):
        """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:2183-2305."""
        pts = np.ascontiguousarray(points, PROJ_POINT_DTYPE)
        d = np.ascontiguousarray(mp_desc, np.uint8).reshape(-1, 32)
        taken = self._u8(kp_taken, F.N)
        match = np.zeros(F.N, np.int32)
        v = F.view()
        n = self.lib.orbm_search_by_projection_reloc(self.ctx, C.byref(v), pts.ctypes.data, d.ctypes.data, len(pts),
                                                     float(th), int(ORBdist), int(self.mbCheckOrientation),
                                                     taken.ctypes.data, match.ctypes.data)
        if n < 0:
            check(n)
        return n, match

    def Fuse(self, KF: Frame, points, mp_desc, th: float = 3.0, inv_level_sigma2=None, idx_offset: int = 0):
        """The search of Fuse (src/ORBmatcher.cc:1403-1737): returns (nFused, best_idx[m], best_dist[m])."""
        pts = np.ascontiguousarray(points, PROJ_POINT_DTYPE)
        d = np.ascontiguousarray(mp_desc, np.uint8).reshape(-1, 32)
        inv = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
        bi = np.zeros(len(pts), np.int32)
        bd = np.zeros(len(pts), np.int32)
        v = KF.view()
        n = self.lib.orbm_fuse(self.ctx, C.byref(v), pts.ctypes.data, d.ctypes.data, len(pts), float(th),
                               None if inv is None else inv.ctypes.data, int(idx_offset), bi.ctypes.data, bd.ctypes.data)
        if n < 0:
            check(n)
        return n, bi, bd

    def SearchBySim3(self, KF1: Frame, KF2: Frame, points1, mp_desc1, points2, mp_desc2, th: float):
        """SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th), src/ORBmatcher.cc:1739-1963.
        Returns (nFound, match12[KF1.N])."""
        p1 = np.ascontiguousarray(points1, PROJ_POINT_DTYPE)
        p2 = np.ascontiguousarray(points2, PROJ_POINT_DTYPE)
        d1 = np.ascontiguousarray(mp_desc1, np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(mp_desc2, np.uint8).reshape(-1, 32)
        assert len(p1) == KF1.N and len(p2) == KF2.N
        m12 = np.zeros(KF1.N, np.int32)
        v1, v2 = KF1.view(), KF2.view()
        n = self.lib.orbm_search_by_sim3(self.ctx, C.byref(v1), C.byref(v2), p1.ctypes.data, d1.ctypes.data,
                                         p2.ctypes.data, d2.ctypes.data, float(th), m12.ctypes.data)
        if n < 0:
            check(n)
        return n, m12

    def SearchForTriangulation(self, KF1: BowSide, has_mp1, KF2: BowSide, has_mp2, stereo1=None, stereo2=None,
                               bOnlyStereo=False, epipole=None, scale_factors2=None, pair_test=None, F12=None,
                               level_sigma2_2=None):
        """SearchForTriangulation, src/ORBmatcher.cc:969-1401 (both overloads).  pair_test(idx1, idx2) -> bool is
        the camera model's epipolarConstrain / matchAndtriangulate.  Returns (nmatches, vMatches12[KF1.N])."""
        h1, h2 = self._u8(has_mp1, len(KF1.keys)), self._u8(has_mp2, len(KF2.keys))
        s1 = None if stereo1 is None else np.ascontiguousarray(stereo1, np.uint8)
        s2 = None if stereo2 is None else np.ascontiguousarray(stereo2, np.uint8)
        ep = None if epipole is None else np.ascontiguousarray(epipole, np.float32)
        sf = None if scale_factors2 is None else np.ascontiguousarray(scale_factors2, np.float32)
        f12 = None if F12 is None else np.ascontiguousarray(F12, np.float32).reshape(9)
        ls2 = None if level_sigma2_2 is None else np.ascontiguousarray(level_sigma2_2, np.float32)
        nlev = max(len(sf) if sf is not None else 0, len(ls2) if ls2 is not None else 0)
        if sf is not None and ls2 is not None:
            assert len(sf) == len(ls2)
        cb = PAIR_TEST(lambda user, a, b: int(bool(pair_test(a, b)))) if pair_test is not None else None
        m12 = np.zeros(len(KF1.keys), np.int32)
        v1, v2 = KF1.view(), KF2.view()
        p = lambda a: None if a is None else a.ctypes.data  # noqa: E731
        n = self.lib.orbm_search_for_triangulation(
            self.ctx, C.byref(v1), h1.ctypes.data, p(s1), C.byref(v2), h2.ctypes.data, p(s2), int(bOnlyStereo), p(ep),
            p(sf), nlev, C.cast(cb, C.c_void_p) if cb is not None else None, None, p(f12), p(ls2),
            int(self.mbCheckOrientation), m12.ctypes.data)
        if n < 0:
            check(n)
        return n, m12

    def ComputeDistinctiveDescriptors(self, desc, group_ptr):
        """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:330-404) for a batch of map points.
        Returns best[g] (row inside group g, -1 for an empty group)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        gp = np.ascontiguousarray(group_ptr, np.int32)
        best = np.zeros(len(gp) - 1, np.int32)
        check(self.lib.orbm_distinctive_descriptors(self.ctx, d.ctypes.data, gp.ctypes.data, len(gp) - 1,
                                                    best.ctypes.data))
        return best

    # ---- DBoW2 transform ---------------------------------------------------------------
    def load_vocabulary(self, voc: Vocabulary):
        """Uploads the vocabulary tree to the matcher's device (orbm_vocab_create); returns an opaque handle."""
        h = C.c_void_p()
        check(self.lib.orbm_vocab_create(self.ctx, len(voc.weight), voc.child_ptr.ctypes.data, voc.child_ids.ctypes.data,
                                         voc.node_desc.ctypes.data, voc.weight.ctypes.data, voc.word_id.ctypes.data,
                                         voc.L, voc.weighting, voc.norm, C.byref(h)))
        return h

    def free_vocabulary(self, h):
        self.lib.orbm_vocab_destroy(h)

    def transform(self, voc_handle, desc, levelsup=4):
        """mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4), src/Frame.cc:729-736.
        Returns (bow_ids, bow_vals, fv_node_ids, fv_ptr, fv_items)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d)
        bi, bv = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        fn, fp, fi = np.zeros(max(n, 1), np.uint32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.uint32)
        nw, nf = C.c_int32(), C.c_int32()
        check(self.lib.orbm_bow_transform(self.ctx, voc_handle, d.ctypes.data, n, int(levelsup), bi.ctypes.data,
                                          bv.ctypes.data, C.byref(nw), fn.ctypes.data, fp.ctypes.data, fi.ctypes.data,
                                          C.byref(nf)))
        return bi[:nw.value].copy(), bv[:nw.value].copy(), fn[:nf.value].copy(), fp[:nf.value + 1].copy(), \
            fi[:fp[nf.value]].copy()
