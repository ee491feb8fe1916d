"""The matcher oracles without a GPU.  Every test runs twice: on kind="ref" -- the reference's OWN
src/ORBmatcher.cc / Frame.cc / MapPoint.cc / DBoW2 compiled verbatim (oracle/_ref/liborbmref.so, driven
through oracle/ref_match_capi.cpp) -- and on kind="port", our restatement oracle/match_oracle.cpp, which
is thereby pinned to the reference: both must reproduce the committed goldens (generated from "ref"),
numpy popcount, the real cv2.BFMatcher and a brute-force window scan, and they must agree with each
other on extra seeded cases of every search."""
import numpy as np
import pytest

from orb_slam3_mac_b200 import synth as S


@pytest.fixture(scope="module", params=["ref", "port"])
def mo(request, oracle_libs):
    if request.param == "ref" and not oracle_libs.have_mref():
        pytest.skip("oracle/_ref/liborbmref.so needs /root/reference to build")
    return oracle_libs.MatchOracle(request.param)


def _rand_frame(n, w, h, seed, nlevels=8):
    from orb_slam3_mac_b200.matcher import Frame
    from orb_slam3_mac_b200._lib import KP_DTYPE
    rng = np.random.default_rng(seed)
    k = np.zeros(n, KP_DTYPE)
    k["x"] = rng.uniform(-5, w + 5, n)  # a few outside the image bounds (undistorted keypoints can be)
    k["y"] = rng.uniform(-5, h + 5, n)
    k["octave"] = rng.integers(0, nlevels, n)
    k["angle"] = rng.uniform(0, 360, n)
    d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    sf = (np.float32(1.2) ** np.arange(nlevels)).astype(np.float32)
    return Frame(k, d, w, h, sf)


def test_descriptor_distance_is_popcount(mo):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b[:20] = a[:20]
    b[20:40] = ~a[20:40]
    for i in range(200):
        assert mo.distance(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())


def test_knn2_matches_cv2_bfmatcher_including_ties(mo):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(1)
    for nq, nt, lowent in ((300, 257, True), (100, 400, False), (5, 1, False), (7, 2, True)):
        if lowent:  # low-entropy descriptors -> many distance ties
            q = (rng.integers(0, 2, (nq, 32), dtype=np.uint8) * 255).astype(np.uint8)
            t = (rng.integers(0, 2, (nt, 32), dtype=np.uint8) * 255).astype(np.uint8)
        else:
            q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
            t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        idx, dist = mo.knn2(q, t)
        ref = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
        assert len(ref) == nq
        for i, ms in enumerate(ref):
            assert len(ms) == min(2, nt)
            for k, m in enumerate(ms):
                assert (idx[i, k], dist[i, k]) == (m.trainIdx, int(m.distance)), (i, k)
            for k in range(len(ms), 2):
                assert idx[i, k] == -1 and dist[i, k] == -1


def test_features_in_area_oracle_vs_bruteforce_and_product(mo):
    """Order = grid columns outer, rows inner, insertion order inside a cell (src/Frame.cc:682-684)."""
    from orb_slam3_mac_b200.matcher import Frame  # noqa: F401
    F = _rand_frame(3000, 752, 480, 3)
    rng = np.random.default_rng(4)
    k = F.keys
    px = np.round((k["x"] - F.bounds[0]) * F.grid_w_inv).astype(np.int64)
    py = np.round((k["y"] - F.bounds[1]) * F.grid_h_inv).astype(np.int64)
    ingrid = (px >= 0) & (px < 64) & (py >= 0) & (py < 48)
    total = 0
    for it in range(300):
        x, y = float(rng.uniform(-30, 780)), float(rng.uniform(-30, 510))
        r = float(rng.choice([3.0, 10.0, 40.0, 100.0]))
        mn, mx = [(-1, -1), (0, 0), (2, 3), (1, -1), (0, 7)][it % 5]
        got = mo.features_in_area(F, x, y, r, mn, mx)
        # brute force over all keypoints: window + level gates, restricted to visited cells
        x32, y32, r32 = np.float32(x), np.float32(y), np.float32(r)
        c0 = max(0, int(np.floor((x32 - F.bounds[0] - r32) * F.grid_w_inv)))
        c1 = min(63, int(np.ceil((x32 - F.bounds[0] + r32) * F.grid_w_inv)))
        r0 = max(0, int(np.floor((y32 - F.bounds[1] - r32) * F.grid_h_inv)))
        r1 = min(47, int(np.ceil((y32 - F.bounds[1] + r32) * F.grid_h_inv)))
        ok = ingrid & (px >= c0) & (px <= c1) & (py >= r0) & (py <= r1)
        ok &= (np.abs(k["x"] - x32) < r32) & (np.abs(k["y"] - y32) < r32)
        if mn > 0 or mx >= 0:
            ok &= k["octave"] >= mn
            if mx >= 0:
                ok &= k["octave"] <= mx
        idx = np.nonzero(ok)[0]
        order = np.lexsort((idx, py[idx], px[idx]))
        assert np.array_equal(got, idx[order].astype(np.int32)), it
        total += len(got)
    assert total > 5000


def test_oracle_matches_committed_goldens(mo, oracle_libs):
    """tests/golden/match_golden.json holds what the reference's own matcher code produced (make_golden_match.py);
    both oracles reproduce it, and the input-stage primitives reproduce what the REAL OpenCV (cv2.remap +
    cv2.cvtColor) produced."""
    import json
    import os
    from tests.golden import match_cases as mc
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "match_golden.json")))
    assert mc.run_matcher_cases(mo) == gold["matcher"]
    bgr, mx, my = mc.input_stage_case()
    g = gold["input_stage"]
    assert mc.sha(oracle_libs.cvt_gray(oracle_libs.remap_linear(bgr, mx, my), False, 15)) == g["bgr_rect_gray"]
    assert mc.sha(oracle_libs.remap_linear(bgr[:, :, 1].copy(), mx, my)) == g["rect_gray_only"]
    assert mc.sha(oracle_libs.cvt_gray(bgr, True, 15)) == g["rgb_gray"]


def test_n3_oracle_matches_committed_goldens(mo):
    """tests/golden/n3_golden.json: SearchByBoW(KF,KF), Sim3 / reloc SearchByProjection, Fuse x2, SearchBySim3,
    SearchForTriangulation x2, ComputeDistinctiveDescriptors and the DBoW2 transform as the reference's own code
    computed them (make_golden_n3.py); both oracles reproduce them."""
    import json
    import os
    from tests.golden import n3_cases as nc
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "n3_golden.json")))
    assert nc.digest(nc.run_n3_cases(mo)) == gold


def test_distinctive_descriptors_oracle_vs_numpy(mo):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:330-404) restated a second time in numpy:
    N x N popcount matrix, row sort, element (int)(0.5 * (N - 1)), first minimum."""
    from tests.golden import n3_cases as nc
    d, gp = nc.distinctive_case(8)
    best = mo.distinctive_descriptors(d, gp)
    for g in range(len(gp) - 1):
        rows = d[gp[g]:gp[g + 1]]
        if len(rows) == 0:
            assert best[g] == -1
            continue
        D = np.unpackbits(rows[:, None, :] ^ rows[None, :, :], axis=2).sum(axis=2)
        med = np.sort(D, axis=1)[:, int(0.5 * (len(rows) - 1))]
        assert best[g] == int(np.argmin(med))
