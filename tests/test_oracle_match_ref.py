"""Pins oracle/match_oracle.cpp (our restatement, kind="port") to the reference's OWN matcher code compiled
verbatim (kind="ref": src/ORBmatcher.cc, src/Frame.cc, src/MapPoint.cc, DBoW2 -> oracle/_ref/liborbmref.so) on
seeded cases beyond the committed goldens: other seeds, window sizes, stereo gates, forward / backward motion,
already-attached keypoints, the fisheye forms, Frame::ComputeStereoMatches on a real extracted stereo pair, and
Frame::UndistortKeyPoints against the real cv2.undistortPoints.  Bar: identical indices / counts / float bits."""
import numpy as np
import pytest

from orb_slam3_mac_b200 import synth as S
from tests.golden import match_cases as mc

SF = mc.SF


@pytest.fixture(scope="module")
def both(oracle_libs):
    if not oracle_libs.have_mref():
        pytest.skip("oracle/_ref/liborbmref.so needs /root/reference to build")
    return oracle_libs.MatchOracle("port"), oracle_libs.MatchOracle("ref")


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if isinstance(x, np.ndarray):
            assert np.array_equal(x, y)
        else:
            assert x == y


@pytest.mark.parametrize("seed", [2, 3])
def test_tracking_searches_port_equals_reference(both, seed):
    from orb_slam3_mac_b200.matcher import Frame, RightHalf
    port, ref = both
    rng = np.random.default_rng(100 + seed)
    (ka, da), (kb, db) = mc.frames(seed, n=1500)
    ur = np.where(rng.random(len(kb)) < 0.6, kb["x"] - rng.uniform(1, 40, len(kb)), -1.0).astype(np.float32)
    FB, FBs = Frame(kb, db, 752, 480, SF), Frame(kb, db, 752, 480, SF, u_right=ur)
    obs0 = np.where(rng.random(len(kb)) < 0.2, rng.integers(0, 3, len(kb)), -1).astype(np.int32)
    # M2: window factor 1 (bFactor false) and 3, far-point gate, stereo gate, keypoints that already hold a point
    mps, md = S.map_points_from_frame(ka, da, 752, 480, 4000, seed)
    for F, th, far in ((FB, 1.0, False), (FBs, 3.0, True), (FBs, 1.0, False)):
        _same(port.search_by_projection_points(F, mps, md, th, far, 30.0, 0.8, kp_obs=obs0),
              ref.search_by_projection_points(F, mps, md, th, far, 30.0, 0.8, kp_obs=obs0))
    # M3: mono, forward, backward; with the stereo gate (mbf != 0) and without the orientation check
    lo, ld = S.last_frame_obs(ka, da, seed)
    for F, th, fwd, bwd, mbf, check in ((FB, 15.0, False, False, 0.0, True), (FBs, 7.0, True, False, 40.0, True),
                                        (FBs, 7.0, False, True, 40.0, False)):
        _same(port.search_by_projection_frames(F, lo, ld, th, fwd, bwd, mbf, check, kp_obs=obs0),
              ref.search_by_projection_frames(F, lo, ld, th, fwd, bwd, mbf, check, kp_obs=obs0))
    # M4: every octave-0 keypoint of F1 against windows of 10 / 100 px
    k0, k1 = ka.copy(), kb.copy()
    k0["octave"] = np.where(rng.random(len(k0)) < 0.7, 0, 2)
    k1["octave"] = np.where(rng.random(len(k1)) < 0.8, 0, 1)
    F0, F1 = Frame(k0, da, 752, 480, SF), Frame(k1, db, 752, 480, SF)
    prev = np.stack([k0["x"] + 3, k0["y"]], axis=1).astype(np.float32)
    for window, ratio, check in ((10, 0.9, True), (100, 0.9, True), (40, 0.6, False)):
        _same(port.search_for_initialization(F0, F1, prev, window, ratio, check),
              ref.search_for_initialization(F0, F1, prev, window, ratio, check))
    # fisheye stereo forms of M2 / M3
    FA = Frame(ka, da, 752, 480, SF)
    R = RightHalf(kb, db, np.where(np.arange(len(ka)) % 3 == 0, np.arange(len(ka)) % len(kb), -1).astype(np.int32),
                  np.where(np.arange(len(kb)) % 4 == 0, np.arange(len(kb)) % len(ka), -1).astype(np.int32))
    uvr = np.stack([lo["u"] - 3, lo["v"]], axis=1).astype(np.float32)
    obs2 = np.where(rng.random(len(ka) + len(kb)) < 0.2, rng.integers(0, 3, len(ka) + len(kb)), -1).astype(np.int32)
    lo2 = lo.copy()
    lo2["u"] -= 3.0
    for th, fwd, bwd, check in ((15.0, False, False, True), (7.0, True, False, False), (7.0, False, True, True)):
        _same(port.search_by_projection_frames_stereo(FA, R, lo2, uvr, ld, th, fwd, bwd, check, kp_obs=obs2),
              ref.search_by_projection_frames_stereo(FA, R, lo2, uvr, ld, th, fwd, bwd, check, kp_obs=obs2))


def test_compute_stereo_matches_port_equals_reference(both, oracle_libs):
    """Frame::ComputeStereoMatches (src/Frame.cc:802-980) compiled verbatim, reading the reference extractor's own
    padded pyramids, against the restatement: mvuRight / mvDepth bit-identical."""
    from orb_slam3_mac_b200.matcher import Frame
    port, ref = both
    ol = oracle_libs.OracleExtractor(1200, 1.2, 8, 20, 7, kind="ref")
    orr = oracle_libs.OracleExtractor(1200, 1.2, 8, 20, 7, kind="ref")
    for seed in (1, 2):
        l, r = S.stereo_pair(480, 752, seed)
        _, kl, dl = ol(l, (0, 0))
        _, kr, dr = orr(r, (0, 0))
        L, R = Frame(kl, dl, 752, 480, SF), Frame(kr, dr, 752, 480, SF)
        pl = [ol.level(i, 19) for i in range(8)]
        pr = [orr.level(i, 19) for i in range(8)]
        for mb, mbf in ((40.0 * 0.11, 435.2 * 0.11), (1.0, 60.0)):
            pu, pd, _ = port.compute_stereo_matches(L, R, pl, pr, mb, mbf)
            ru, rd, _ = ref.compute_stereo_matches(L, R, pl, pr, mb, mbf)
            assert np.array_equal(pu.view(np.uint32), ru.view(np.uint32)), (seed, mb)
            assert np.array_equal(pd.view(np.uint32), rd.view(np.uint32)), (seed, mb)
        assert (ru >= 0).sum() > 100   # (the last baseline admits the pair's 5..60 px disparities)


def test_reference_undistort_keypoints_vs_cv2(oracle_libs):
    """Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:738-800) compiled verbatim on the shim's
    cv::undistortPoints, against the real cv2.undistortPoints with the same K, distortion and P = K."""
    import ctypes as C
    cv2 = pytest.importorskip("cv2")
    if not oracle_libs.have_mref():
        pytest.skip("needs oracle/_ref/liborbmref.so")
    lib = C.CDLL(oracle_libs.MREF_SO)
    from orb_slam3_mac_b200._lib import KP_DTYPE
    rng = np.random.default_rng(5)
    for (fx, fy, cx, cy), dist, (w, h) in (((458.654, 457.296, 367.215, 248.375), [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], (752, 480)),
                                           ((535.4, 539.2, 320.1, 247.6), [0.2624, -0.9531, -0.0054, 0.0026, 1.1633], (640, 480)),
                                           ((500.0, 500.0, 320.0, 240.0), [0.0, 0.0, 0.0, 0.0], (640, 480))):
        n = 2000
        keys = np.zeros(n, KP_DTYPE)
        keys["x"] = rng.uniform(0, w, n).astype(np.float32)
        keys["y"] = rng.uniform(0, h, n).astype(np.float32)
        d = np.array(dist, np.float32)
        out = np.zeros((n, 2), np.float32)
        bounds = np.zeros(4, np.float32)
        rc = lib.orbr_undistort_keypoints(C.c_void_p(keys.ctypes.data), n, C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy),
                                          C.c_void_p(d.ctypes.data), len(d), w, h, C.c_void_p(out.ctypes.data),
                                          C.c_void_p(bounds.ctypes.data))
        assert rc == 0
        K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
        if d[0] == 0.0:   # src/Frame.cc:740-744: no distortion, keypoints and bounds are the image's
            assert np.array_equal(out[:, 0], keys["x"]) and np.array_equal(out[:, 1], keys["y"])
            assert list(bounds) == [0.0, float(w), 0.0, float(h)]
            continue
        pts = np.stack([keys["x"], keys["y"]], axis=1).reshape(-1, 1, 2).astype(np.float32)
        want = cv2.undistortPoints(pts, K, d, None, K).reshape(-1, 2)
        assert np.array_equal(out.view(np.uint32), want.view(np.uint32))
        corners = np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32).reshape(-1, 1, 2)
        c = cv2.undistortPoints(corners, K, d, None, K).reshape(-1, 2)
        assert list(bounds) == [min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0]), min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1])]
