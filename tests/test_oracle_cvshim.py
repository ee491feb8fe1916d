"""oracle/cvshim's restatements of the OpenCV calls that the verbatim matcher build (oracle/_ref/liborbmref.so)
reaches, against the REAL OpenCV of this image (cv2): cv::gemm (what every `A * B`, `A * p + t`, `-R.t() * t` of
src/ORBmatcher.cc / src/Frame.cc becomes), Mat::inv() of a 3x3, cv::norm, cv::undistortPoints and
BFMatcher(NORM_HAMMING).knnMatch.  Bar: identical float bits."""
import ctypes as C

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


@pytest.fixture(scope="module")
def lib(oracle_libs):
    if not oracle_libs.have_mref():
        pytest.skip("oracle/_ref/liborbmref.so needs /root/reference to build")
    l = C.CDLL(oracle_libs.MREF_SO)
    l.orbr_shim_norm.restype = C.c_double
    return l


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _rot(rng):
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    return np.ascontiguousarray(q, np.float32)


def _gemm(lib, A, B, alpha, Cm, beta, flags):
    m = A.shape[1] if flags & 1 else A.shape[0]
    n = B.shape[0] if flags & 2 else B.shape[1]
    out = np.zeros((m, n), np.float32)
    lib.orbr_shim_gemm(_p(A), A.shape[0], A.shape[1], _p(B), B.shape[0], B.shape[1], C.c_double(alpha), _p(Cm),
                       C.c_double(beta), flags, _p(out))
    return out


def test_gemm_matches_cv2_on_pose_shapes(lib):
    rng = np.random.default_rng(0)
    bad = 0
    for it in range(300):
        for (ar, ac, bc) in ((3, 3, 1), (3, 3, 3), (3, 4, 1), (4, 4, 4), (2, 3, 3)):
            for flags in (0, 1, 2):
                A = rng.normal(size=(ar, ac)).astype(np.float32) * np.float32(10 ** rng.uniform(-2, 3))
                sh_b = (ac, bc) if not flags & 2 else (bc, ac)
                if flags & 1:
                    A = np.ascontiguousarray(A.T)
                B = rng.normal(size=sh_b).astype(np.float32)
                m = A.shape[1] if flags & 1 else A.shape[0]
                n = B.shape[0] if flags & 2 else B.shape[1]
                for alpha, beta in ((1.0, 0.0), (-1.0, 0.0), (1.0, 1.0), (-1.0, 1.0), (1.0, -1.0), (2.0, 0.0)):   # the forms in use
                    Cm = rng.normal(size=(m, n)).astype(np.float32) if beta else None
                    got = _gemm(lib, A, B, alpha, Cm, beta, flags)
                    want = cv2.gemm(A, B, alpha, Cm, beta, flags=flags)
                    bad += int(not np.array_equal(got.view(np.uint32), want.view(np.uint32)))
    assert bad == 0


def test_pose_expressions_match_cv2(lib):
    """The MatExpr forms of the compiled reference files, evaluated as OpenCV folds them into gemm calls."""
    rng = np.random.default_rng(1)
    for it in range(200):
        R1, R2 = _rot(rng), _rot(rng)
        p = rng.normal(size=(3, 1)).astype(np.float32) * 5
        t = rng.normal(size=(3, 1)).astype(np.float32)
        s = np.float32(rng.uniform(0.5, 2.0))
        out = np.zeros(9, np.float32)
        exp = {0: cv2.gemm(R1, p, 1, t, 1),                                   # Rcw * p3Dw + tcw
               1: cv2.gemm(R1, t, -1, None, 0, flags=cv2.GEMM_1_T),           # -Rcw.t() * tcw
               2: cv2.gemm(R1, R2, 1, None, 0, flags=cv2.GEMM_2_T),           # R1w * R2w.t()
               3: (R1 * np.float32(1.0 / np.float64(s))).astype(np.float32)}  # sRcw / scw
        exp[5] = cv2.gemm(cv2.gemm(R1, R2, -1, None, 0, flags=cv2.GEMM_2_T), p, 1, t, 1)   # -R1w * R2w.t() * t2w + t1w
        for which, want in exp.items():
            n = lib.orbr_shim_expr(which, _p(R1), _p(R2), _p(p), _p(t), C.c_float(s), _p(out))
            assert n == want.size
            assert np.array_equal(out[:n].view(np.uint32), want.reshape(-1).view(np.uint32)), (it, which)


def test_invert3_and_norm_match_cv2(lib):
    rng = np.random.default_rng(2)
    for it in range(300):
        M = (np.eye(3) * rng.uniform(100, 800) + rng.normal(size=(3, 3)) * rng.uniform(0, 50)).astype(np.float32)
        out = np.zeros((3, 3), np.float32)
        lib.orbr_shim_invert3(_p(M), _p(out))
        ok, want = cv2.invert(M)
        assert np.array_equal(out.view(np.uint32), want.view(np.uint32)), it
        v = (rng.normal(size=(3, 1)) * 10 ** rng.uniform(-2, 3)).astype(np.float32)
        assert lib.orbr_shim_norm(_p(v), 3) == cv2.norm(v)


def test_undistort_points_matches_cv2(lib):
    rng = np.random.default_rng(3)
    for dist in ([-0.2834, 0.0739, 0.00019, 1.76e-05], [0.2624, -0.9531, -0.0054, 0.0026, 1.1633], [-0.4, 0.25, 0.001, -0.002, -0.09]):
        K = np.array([[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]], np.float32)
        P = np.array([[435.2, 0, 367.4], [0, 435.2, 252.2], [0, 0, 1]], np.float32)
        d = np.array(dist, np.float32)
        n = 3000
        xy = np.stack([rng.uniform(-20, 780, n), rng.uniform(-20, 500, n)], axis=1).astype(np.float32)
        out = np.zeros((n, 2), np.float32)
        lib.orbr_shim_undistort(_p(xy), n, _p(K), _p(d), len(d), _p(P), _p(out))
        want = cv2.undistortPoints(xy.reshape(-1, 1, 2), K, d, None, P).reshape(-1, 2)
        assert np.array_equal(out.view(np.uint32), want.view(np.uint32))


def test_bfmatcher_knn_matches_cv2_including_ties(lib):
    rng = np.random.default_rng(4)
    for nq, nt, lowent in ((300, 257, True), (100, 400, False), (5, 1, False)):
        mk = (lambda n: (rng.integers(0, 2, (n, 32), dtype=np.uint8) * 255).astype(np.uint8)) if lowent else \
             (lambda n: rng.integers(0, 256, (n, 32), dtype=np.uint8))
        q, t = mk(nq), mk(nt)
        idx = np.zeros((nq, 2), np.int32)
        dist = np.zeros((nq, 2), np.float32)
        lib.orbr_shim_knn2(_p(q), nq, _p(t), nt, _p(idx), _p(dist))
        ref = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
        for i, ms in enumerate(ref):
            for k, m in enumerate(ms):
                assert (idx[i, k], dist[i, k]) == (m.trainIdx, m.distance)
            for k in range(len(ms), 2):
                assert idx[i, k] == -1
