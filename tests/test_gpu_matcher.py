"""Parity of the CUDA Hamming searches (through the C ABI of include/orbm.h) on the same POD snapshots against
the REFERENCE'S OWN matcher code -- src/ORBmatcher.cc, src/Frame.cc, src/MapPoint.cc, DBoW2 compiled verbatim into
oracle/_ref/liborbmref.so (MatchOracle("ref"); the prebuilt library travels to the GPU box) -- and, for the one
piece the reference has no separate function for (the row search inside ComputeStereoMatches), against our
restatement, which tests/test_oracle_match*.py pin to the reference.  Bar: bit-exact indices and counts."""
import os

import numpy as np
import pytest

from orb_slam3_mac_b200 import synth as S

pytestmark = pytest.mark.gpu

SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


@pytest.fixture(scope="module")
def mo(oracle_libs):
    return oracle_libs.MatchOracle("ref" if oracle_libs.have_mref() else "port")


@pytest.fixture(scope="module")
def mo_port(oracle_libs):
    return oracle_libs.MatchOracle("port")


@pytest.fixture(scope="module")
def matcher():
    from orb_slam3_mac_b200.matcher import ORBmatcher
    m = ORBmatcher(0.8, True)
    yield m
    m.close()


@pytest.fixture(scope="module")
def extract(oracle_libs):
    """Frames for the matcher tests come from the CPU oracle extractor (any extractor would do:
    both sides of each comparison see identical snapshots)."""
    cache = {}

    def run(img, nf, lap=(0, 0)):
        key = (hash(img.tobytes()), img.shape, nf, lap)
        if key not in cache:
            kind = "ref" if oracle_libs.have_ref() else "port"
            ex = oracle_libs.OracleExtractor(nf, 1.2, 8, 20, 7, kind=kind)
            cache[key] = ex(img, lap)
        return cache[key]

    return run


def _frame(keys, desc, w, h, u_right=None):
    from orb_slam3_mac_b200.matcher import Frame
    return Frame(keys, desc, w, h, SF, u_right=u_right)


def test_hamming_csr_matches_popcount(matcher):
    rng = np.random.default_rng(0)
    q = rng.integers(0, 256, (700, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (1300, 32), dtype=np.uint8)
    counts = rng.integers(0, 40, 700)
    counts[::7] = 0  # ragged: empty rows
    row_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
    cand = rng.integers(0, 1300, row_ptr[-1]).astype(np.int32)
    got = matcher.hamming_csr(q, t, row_ptr, cand)
    rows = np.repeat(np.arange(700), counts)
    exp = np.unpackbits(q[rows] ^ t[cand], axis=1).sum(axis=1)
    assert np.array_equal(got, exp.astype(np.uint16))
    assert len(matcher.hamming_csr(q, t, np.zeros(701, np.int32), np.zeros(0, np.int32))) == 0


def test_knn2_matches_oracle_and_cv2(matcher, mo):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(2)
    for nq, nt, lowent in ((1500, 1500, False), (300, 257, True), (64, 1, False), (10, 2, True), (33, 1564, False)):
        if lowent:
            q = (rng.integers(0, 2, (nq, 32), dtype=np.uint8) * 255).astype(np.uint8)
            t = (rng.integers(0, 2, (nt, 32), dtype=np.uint8) * 255).astype(np.uint8)
        else:
            q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
            t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        idx, dist = matcher.knn2(q, t)
        oi, od = mo.knn2(q, t)
        assert np.array_equal(idx, oi) and np.array_equal(dist, od)
        ref = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q[:50], t, k=2)
        for i, ms in enumerate(ref):
            for k, m in enumerate(ms):
                assert (idx[i, k], dist[i, k]) == (m.trainIdx, int(m.distance))


def test_fisheye_stereo_knn2_c3(matcher, mo, extract):
    """Config C3: 2 x 512x512, 1500 features, lapping {0,511} -> every keypoint is 'stereo'
    (monoLeft = monoRight = 0), brute-force 2-NN + Lowe ratio 0.7 (src/Frame.cc:1146-1153)."""
    l, r = S.fisheye_pair(512, 5)
    ml, kl, dl = extract(l, 1500, (0, 511))
    mr, kr, dr = extract(r, 1500, (0, 511))
    assert ml == 0 and mr == 0
    idx, dist = matcher.knn2(dl, dr)
    oi, od = mo.knn2(dl, dr)
    assert np.array_equal(idx, oi) and np.array_equal(dist, od)
    good = dist[:, 0] < dist[:, 1] * np.float32(0.7)
    assert good.sum() > 100  # the pair really overlaps


def test_stereo_rows_c2(matcher, mo_port, extract):
    """Config C2: EuRoC stereo pair, 1200 features per image, lapping {0,0}."""
    for seed in (1, 2):
        l, r = S.stereo_pair(480, 752, seed)
        _, kl, dl = extract(l, 1200)
        _, kr, dr = extract(r, 1200)
        L, R = _frame(kl, dl, 752, 480), _frame(kr, dr, 752, 480)
        for min_d, max_d in ((0.0, 80.0), (0.0, 20.0), (5.0, 400.0)):
            bi, bd = matcher.stereo_rows(L, R, min_d, max_d)
            oi, od = mo_port.stereo_rows(L, R, 480, min_d, max_d)
            assert np.array_equal(bi, oi) and np.array_equal(bd, od)
        bi, bd = matcher.stereo_rows(L, R, 0.0, 80.0)
        assert (bd < 75).sum() > 200  # thOrbDist (src/Frame.cc:807): plenty of real matches


@pytest.mark.parametrize("th,far,u_right", [(1.0, False, False), (3.0, True, False), (1.0, False, True)])
def test_search_by_projection_points_c4(matcher, mo, extract, th, far, u_right):
    """Config C4: 1280x720, 2000 features, 10,000 map points."""
    img0 = S.textured_noise(720, 1280, 7)
    img1 = np.roll(img0, 3, axis=1)  # same scene, 3-px shift
    _, k0, d0 = extract(img0, 2000, (0, 1000))
    _, k1, d1 = extract(img1, 2000, (0, 1000))
    mps, md = S.map_points_from_frame(k0, d0, 1280, 720, 10000, 7)
    rng = np.random.default_rng(3)
    ur = None
    if u_right:
        ur = np.where(rng.random(len(k1)) < 0.6, k1["x"] - rng.uniform(1, 40, len(k1)), -1.0).astype(np.float32)
    F = _frame(k1, d1, 1280, 720, u_right=ur)
    obs0 = np.where(rng.random(len(k1)) < 0.2, rng.integers(0, 3, len(k1)), -1).astype(np.int32)
    n, match, obs = matcher.SearchByProjection(F, mps, md, th, far, 30.0, kp_obs=obs0)
    on, omatch, oobs = mo.search_by_projection_points(F, mps, md, th, far, 30.0, 0.8, kp_obs=obs0)
    assert n == on and np.array_equal(match, omatch) and np.array_equal(obs, oobs)
    assert n > 300


def test_window_searches_of_changing_size_on_one_context(mo, extract):
    """The window searches size their candidate lists from the previous search of the same context and fetch them on
    that guess: a context that sees small, 10x larger, small again, then a 100-px-window search and back must return
    the oracle's result every time (first call and growth: the fill reruns; moderate growth: the tail is fetched
    after the count; shrink: the guess over-fetches)."""
    from orb_slam3_mac_b200.matcher import ORBmatcher
    img0 = S.textured_noise(720, 1280, 9)
    img1 = np.roll(img0, 2, axis=1)
    _, k0, d0 = extract(img0, 2000, (0, 1000))
    _, k1, d1 = extract(img1, 2000, (0, 1000))
    mps, md = S.map_points_from_frame(k0, d0, 1280, 720, 6000, 9)
    F0, F = _frame(k0, d0, 1280, 720), _frame(k1, d1, 1280, 720)
    prev = np.stack([k0["x"], k0["y"]], axis=1)
    m = ORBmatcher(0.8, True)
    for th in (1.0, 6.0, 1.0, None, 2.0, 1.3, 12.0, 1.0):
        if th is None:
            n, m12, pv = m.SearchForInitialization(F0, F, prev, 100)
            on, om12, opv = mo.search_for_initialization(F0, F, prev, 100, 0.8, True)
            assert n == on and np.array_equal(m12, om12) and np.array_equal(pv, opv)
        else:
            n, match, obs = m.SearchByProjection(F, mps, md, th, False, 30.0)
            on, omatch, oobs = mo.search_by_projection_points(F, mps, md, th, False, 30.0, 0.8)
            assert n == on and np.array_equal(match, omatch) and np.array_equal(obs, oobs), th
    m.close()


@pytest.mark.parametrize("fwd,bwd,check,u_right", [(False, False, True, False), (True, False, True, False),
                                                    (False, True, False, True)])
def test_search_by_projection_frames(matcher, mo, extract, fwd, bwd, check, u_right):
    from orb_slam3_mac_b200.matcher import ORBmatcher
    img0 = S.textured_noise(480, 752, 11)
    img1 = np.roll(img0, 3, axis=1)
    _, k0, d0 = extract(img0, 1000, (0, 1000))
    _, k1, d1 = extract(img1, 1000, (0, 1000))
    lo, ld = S.last_frame_obs(k0, d0, 11)
    rng = np.random.default_rng(5)
    ur = (k1["x"] - rng.uniform(1, 30, len(k1))).astype(np.float32) if u_right else None
    cur = _frame(k1, d1, 752, 480, u_right=ur)
    m = ORBmatcher(0.9, check)
    for th in (7.0, 15.0):
        n, match, obs = m.SearchByProjectionFrames(cur, lo, ld, th, fwd, bwd, 40.0)
        on, omatch, oobs = mo.search_by_projection_frames(cur, lo, ld, th, fwd, bwd, 40.0, check)
        assert n == on and np.array_equal(match, omatch) and np.array_equal(obs, oobs)
    assert n > 100
    m.close()


def test_search_for_initialization(mo, extract):
    """Mono initialisation uses the 5 x nFeatures extractor (src/Tracking.cc:212) and 100-px
    windows (src/Tracking.cc:1506-1507): hundreds of candidates per query."""
    from orb_slam3_mac_b200.matcher import ORBmatcher
    img0 = S.textured_noise(480, 752, 21)
    img1 = np.roll(img0, 5, axis=1)
    img1 = np.roll(img1, 2, axis=0)
    _, k0, d0 = extract(img0, 5000, (0, 1000))
    _, k1, d1 = extract(img1, 5000, (0, 1000))
    F1, F2 = _frame(k0, d0, 752, 480), _frame(k1, d1, 752, 480)
    prev = np.stack([k0["x"], k0["y"]], axis=1)
    for ratio, check, win in ((0.9, True, 100), (0.9, False, 30)):
        m = ORBmatcher(ratio, check)
        n, m12, pv = m.SearchForInitialization(F1, F2, prev, win)
        on, om12, opv = mo.search_for_initialization(F1, F2, prev, win, ratio, check)
        assert n == on and np.array_equal(m12, om12) and np.array_equal(pv, opv)
        assert n > 100
        # second call with the updated vbPrevMatched, as Tracking does frame after frame
        n2, m12b, _ = m.SearchForInitialization(F1, F2, pv, win)
        on2, om12b, _ = mo.search_for_initialization(F1, F2, opv, win, ratio, check)
        assert n2 == on2 and np.array_equal(m12b, om12b)
        m.close()


@pytest.mark.parametrize("n_a,n_b,step", [(12, 10, 1), (9, 60, 2), (5, 3, 4), (20, 40, 0)])
def test_search_for_initialization_contested_window(mo, n_a, n_b, step):
    """SearchForInitialization when the frame-1 keypoints of one spot all prefer the same frame-2 keypoints: later
    rows find their nearest candidates already matched at a smaller distance (more of them than the keys
    k_csr_topk returns), so they take the full scan of the window; rows shorter than the key count and a window of
    60 candidates (two warp rounds) included; step 0 = every distance equal."""
    from orb_slam3_mac_b200._lib import KP_DTYPE
    from orb_slam3_mac_b200.matcher import ORBmatcher
    A, B = _contested_node_case(n_a, n_b, step, 5 * n_a + step)
    rng = np.random.default_rng(n_a)
    ka, kb = np.zeros(n_a, KP_DTYPE), np.zeros(n_b, KP_DTYPE)
    ka["x"], ka["y"] = 300 + rng.uniform(-3, 3, n_a), 200 + rng.uniform(-3, 3, n_a)
    kb["x"], kb["y"] = 300 + rng.uniform(-40, 40, n_b), 200 + rng.uniform(-40, 40, n_b)
    ka["angle"], kb["angle"] = rng.uniform(0, 359, n_a), rng.uniform(0, 359, n_b)
    F1, F2 = _frame(ka, A.desc, 752, 480), _frame(kb, B.desc, 752, 480)
    prev = np.stack([ka["x"], ka["y"]], axis=1)
    for check in (False, True):
        m = ORBmatcher(0.8, check)
        n, m12, pv = m.SearchForInitialization(F1, F2, prev, 100)
        m.close()
        on, om12, opv = mo.search_for_initialization(F1, F2, prev, 100, 0.8, check)
        assert n == on and np.array_equal(m12, om12) and np.array_equal(pv, opv)
    if step in (1, 2):
        assert on >= 0 and (om12 >= 0).sum() == on


def test_matcher_contexts_are_concurrent(mo):
    """Tracking, LocalMapping and LoopClosing construct ORBmatcher objects on three threads."""
    import threading
    from orb_slam3_mac_b200.matcher import ORBmatcher
    rng = np.random.default_rng(9)
    q = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (900, 32), dtype=np.uint8)
    exp = mo.knn2(q, t)
    res = [None] * 3

    def work(i):
        m = ORBmatcher(0.8, True)
        for _ in range(10):
            res[i] = m.knn2(q, t)
        m.close()

    ts = [threading.Thread(target=work, args=(i,)) for i in range(3)]
    [x.start() for x in ts]
    [x.join() for x in ts]
    for r in res:
        assert np.array_equal(r[0], exp[0]) and np.array_equal(r[1], exp[1])


def test_compute_stereo_matches_full_c2(matcher, mo, oracle_libs):
    """the whole Frame::ComputeStereoMatches (Hamming rows + 11x11 SAD over +-5 px
    on the two DEVICE-RESIDENT pyramids + parabola + median cut) against the CPU restatement that
    reads the reference's padded pyramids.  Bar: mvuRight / mvDepth bit-identical floats."""
    from orb_slam3_mac_b200.extractor import ORBextractor
    kind = "ref" if oracle_libs.have_ref() else "port"
    exl = ORBextractor(1200, 1.2, 8, 20, 7, max_size=(480, 752), max_batch=1)
    exr = ORBextractor(1200, 1.2, 8, 20, 7, max_size=(480, 752), max_batch=1)
    ol = oracle_libs.OracleExtractor(1200, 1.2, 8, 20, 7, kind=kind)
    orr = oracle_libs.OracleExtractor(1200, 1.2, 8, 20, 7, kind=kind)
    try:
        for seed in (1, 2, 3):
            l, r = S.stereo_pair(480, 752, seed)
            _, kl, dl = exl(l, (0, 0))
            _, kr, dr = exr(r, (0, 0))
            ol(l, (0, 0))
            orr(r, (0, 0))
            L, R = _frame(kl, dl, 752, 480), _frame(kr, dr, 752, 480)
            pl = [ol.level(i, 19) for i in range(8)]
            pr = [orr.level(i, 19) for i in range(8)]
            for mb, mbf in ((40.0 * 0.11, 435.2 * 0.11 * 1.0), (1.0, 60.0), (0.5, 200.0)):
                u, d = matcher.ComputeStereoMatches(L, R, exl.device_pyramid(), exr.device_pyramid(), mb, mbf)
                ou, od, nacc = mo.compute_stereo_matches(L, R, pl, pr, mb, mbf)
                assert np.array_equal(u.view(np.uint32), ou.view(np.uint32)), (seed, mb, mbf)
                assert np.array_equal(d.view(np.uint32), od.view(np.uint32)), (seed, mb, mbf)
            assert (u >= 0).sum() > 100  # the synthetic pair has real row-aligned matches
    finally:
        exl.close()
        exr.close()


def test_device_grid_candidates_match_oracle_order(matcher, mo):
    """N2: the device-built 64x48 grid and window queries return exactly the reference's candidate
    lists (order = grid columns outer, rows inner, insertion order inside a cell, src/Frame.cc:682-684)
    and the Hamming distance of every (query, candidate) pair; includes keypoints outside the grid,
    windows that leave the image, invalid queries and every level-gate form."""
    from orb_slam3_mac_b200._lib_match import AREA_QUERY_DTYPE
    rng = np.random.default_rng(11)
    n = 5000
    from orb_slam3_mac_b200._lib import KP_DTYPE
    keys = np.zeros(n, KP_DTYPE)
    keys["x"] = rng.uniform(-20, 772, n).astype(np.float32)     # some land outside the grid
    keys["y"] = rng.uniform(-20, 500, n).astype(np.float32)
    keys["x"][:400] = np.round(keys["x"][:400])                 # crowded integer positions
    keys["y"][:400] = np.round(keys["y"][:400] / 8) * 8
    keys["octave"] = rng.integers(0, 8, n)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    F = _frame(keys, desc, 752, 480)
    nq = 1200
    q = np.zeros(nq, AREA_QUERY_DTYPE)
    q["x"] = rng.uniform(-40, 800, nq).astype(np.float32)
    q["y"] = rng.uniform(-40, 520, nq).astype(np.float32)
    q["r"] = rng.choice(np.array([0.0, 2.5, 7.0, 24.0, 100.0, 900.0], np.float32), nq)
    gates = [(-1, -1), (0, 0), (2, 3), (1, -1), (0, 7), (5, 2)]
    for i in range(nq):
        q["min_level"][i], q["max_level"][i] = gates[i % len(gates)]
    q["valid"] = (rng.uniform(size=nq) > 0.1).astype(np.int32)
    qdesc = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    q["desc_row"] = rng.integers(0, 300, nq)
    row_ptr, cand, dist = matcher.area_candidates(F, q, qdesc, cap=2_000_000)
    assert row_ptr[0] == 0 and row_ptr[-1] == len(cand) and len(cand) > 20000
    pop = np.array([bin(v).count("1") for v in range(256)], np.int32)
    for i in range(nq):
        got = cand[row_ptr[i]:row_ptr[i + 1]]
        if not q["valid"][i]:
            assert len(got) == 0
            continue
        want = mo.features_in_area(F, float(q["x"][i]), float(q["y"][i]), float(q["r"][i]),
                                   int(q["min_level"][i]), int(q["max_level"][i]))
        assert np.array_equal(got, want), i
        if len(got):
            d = pop[desc[got] ^ qdesc[q["desc_row"][i]]].sum(axis=1)
            assert np.array_equal(dist[row_ptr[i]:row_ptr[i + 1]], d.astype(np.uint16)), i
    # without descriptors only the lists come back; too small a capacity is an error, not a truncation
    rp2, cand2, d2 = matcher.area_candidates(F, q, None, cap=2_000_000)
    assert d2 is None and np.array_equal(rp2, row_ptr) and np.array_equal(cand2, cand)
    with pytest.raises(Exception):
        matcher.area_candidates(F, q, None, cap=10)
    # empty frame / no queries
    F0 = _frame(np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8), 752, 480)
    rp0, c0, _ = matcher.area_candidates(F0, q[:5], None, cap=16)
    assert not rp0.any() and len(c0) == 0


def _fisheye_frame(extract, seed):
    """A fisheye stereo frame (config C3): left/right extractions of a 512x512 pair and mutual
    best-match tables standing in for mvLeftToRightMatch / mvRightToLeftMatch."""
    from orb_slam3_mac_b200.matcher import RightHalf
    imgL, imgR = S.fisheye_pair(512, seed)
    _, kl, dl = extract(imgL, 1500, (0, 511))
    _, kr, dr = extract(imgR, 1500, (0, 511))
    rng = np.random.default_rng(seed)
    n = min(len(kl), len(kr))
    pairs = rng.permutation(n)[: n // 3]
    l2r = np.full(len(kl), -1, np.int32)
    r2l = np.full(len(kr), -1, np.int32)
    tgt = rng.permutation(len(kr))[: len(pairs)]
    l2r[pairs] = tgt
    r2l[tgt] = pairs
    return _frame(kl, dl, 512, 512), RightHalf(kr, dr, l2r, r2l), (kl, dl, kr, dr)


@pytest.mark.parametrize("th,far,tables", [(1.0, False, True), (3.0, True, True), (1.0, False, False)])
def test_search_by_projection_points_fisheye_stereo(matcher, mo, extract, th, far, tables):
    """M2 with Frame::Nleft != -1 (src/ORBmatcher.cc:48-218): left branch mirrored through
    mvLeftToRightMatch, right branch in mGridRight mirrored through mvRightToLeftMatch, one shared
    mvpMapPoints state over Nleft + Nright."""
    from orb_slam3_mac_b200._lib_match import MAP_POINT_RIGHT_DTYPE
    from orb_slam3_mac_b200.matcher import RightHalf
    F, R, (kl, dl, kr, dr) = _fisheye_frame(extract, 5)
    if not tables:
        R = RightHalf(kr, dr, None, None)
    m = 6000
    mps, md = S.map_points_from_frame(kl, dl, 512, 512, m, 11)
    mr, mdr = S.map_points_from_frame(kr, dr, 512, 512, m, 12)
    rng = np.random.default_rng(13)
    # every other map point carries the right camera's descriptor, so that both branches find matches
    use_r = rng.random(m) < 0.5
    md = np.where(use_r[:, None], mdr, md)
    mpr = np.zeros(m, MAP_POINT_RIGHT_DTYPE)
    mpr["proj_x"], mpr["proj_y"], mpr["view_cos"] = mr["proj_x"], mr["proj_y"], mr["view_cos"]
    mpr["level"] = np.where(rng.random(m) < 0.05, -1, mr["level"])
    mpr["in_view"] = (rng.random(m) < 0.7).astype(np.uint8)
    mps["in_view"] = (rng.random(m) < 0.7).astype(np.uint8)
    obs0 = np.where(rng.random(F.N + R.N) < 0.2, rng.integers(0, 3, F.N + R.N), -1).astype(np.int32)
    n, match, obs = matcher.SearchByProjectionStereo(F, R, mps, mpr, md, th, far, 30.0, kp_obs=obs0)
    on, omatch, oobs = mo.search_by_projection_points_stereo(F, R, mps, mpr, md, th, far, 30.0, 0.8, kp_obs=obs0)
    assert n == on and np.array_equal(match, omatch) and np.array_equal(obs, oobs)
    assert (match[: F.N] >= 0).sum() > 100 and (match[F.N:] >= 0).sum() > 100


@pytest.mark.parametrize("fwd,bwd,check", [(False, False, True), (True, False, True), (False, True, False)])
def test_search_by_projection_frames_fisheye_stereo(mo, extract, fwd, bwd, check):
    """M3 with CurrentFrame.Nleft != -1 (src/ORBmatcher.cc:1965-2181): the right camera is searched with
    the projection of mTrl * x3Dc after the left one, rotation histogram over both halves."""
    from orb_slam3_mac_b200.matcher import ORBmatcher
    F, R, (kl, dl, kr, dr) = _fisheye_frame(extract, 6)
    lo, ld = S.last_frame_obs(kl, dl, 21)
    rng = np.random.default_rng(22)
    # right-camera projections: near a right keypoint for half of the observations, anywhere otherwise
    j = rng.integers(0, len(kr), len(lo))
    uvr = np.stack([kr["x"][j] + rng.normal(0, 1.0, len(lo)), kr["y"][j] + rng.normal(0, 1.0, len(lo))], axis=1)
    far = rng.random(len(lo)) < 0.5
    uvr[far] = rng.uniform(-50, 560, (int(far.sum()), 2))
    uvr = uvr.astype(np.float32)
    ld = np.where((rng.random(len(lo)) < 0.4)[:, None], S.flip_bits(dr[j], 0.03, rng), ld)
    lo["octave"] = np.where(far, lo["octave"], kr["octave"][j])
    obs0 = np.where(rng.random(F.N + R.N) < 0.15, rng.integers(0, 3, F.N + R.N), -1).astype(np.int32)
    m = ORBmatcher(0.9, check)
    n, match, obs = m.SearchByProjectionFramesStereo(F, R, lo, uvr, ld, 7.0, fwd, bwd, kp_obs=obs0)
    on, omatch, oobs = mo.search_by_projection_frames_stereo(F, R, lo, uvr, ld, 7.0, fwd, bwd, check, kp_obs=obs0)
    m.close()
    assert n == on and np.array_equal(match, omatch) and np.array_equal(obs, oobs)
    assert (match[: F.N] >= 0).sum() > 100 and (match[F.N:] >= 0).sum() > 5


def _feat_vec(keys, desc, rng, nnodes=120, drop=0.0):
    """A stand-in for DBoW2's FeatureVector: node = coarse image cell of the keypoint (matching
    features of two nearby frames mostly share a node), a few features moved to random nodes."""
    node = (np.floor(keys["x"] / 64).astype(np.int64) * 16 + np.floor(keys["y"] / 64).astype(np.int64)) % nnodes
    stray = rng.random(len(keys)) < 0.1
    node[stray] = rng.integers(0, nnodes + 40, int(stray.sum()))
    fv = {}
    for i in rng.permutation(len(keys)):
        if rng.random() < drop:
            continue
        fv.setdefault(int(node[i]) * 7 + 3, []).append(int(i))
    return fv


@pytest.mark.parametrize("check,nleft", [(True, -1), (False, -1), (True, 1)])
def test_search_by_bow_keyframe_frame(mo, extract, check, nleft):
    """N3: SearchByBoW(KeyFrame*, Frame&) (src/ORBmatcher.cc:273-475), mono and fisheye (Nleft != -1)
    forms, node lists that only partly overlap (exercises the lower_bound walk)."""
    from orb_slam3_mac_b200.matcher import BowSide, ORBmatcher
    img0 = S.textured_noise(480, 752, 17)
    img1 = np.roll(img0, 3, axis=1)
    _, k0, d0 = extract(img0, 1500, (0, 0))
    _, k1, d1 = extract(img1, 1500, (0, 0))
    rng = np.random.default_rng(5 + nleft)
    KF = BowSide(k0, d0, _feat_vec(k0, d0, rng, drop=0.05))
    F = BowSide(k1, d1, _feat_vec(k1, d1, rng, drop=0.05))
    ok = (rng.random(len(k0)) < 0.7).astype(np.uint8)
    nl = -1 if nleft < 0 else len(k1) // 2
    m = ORBmatcher(0.7, check)
    n, matches = m.SearchByBoW(KF, ok, F, nl)
    on, omatches = mo.search_by_bow(KF, ok, F, nl, 0.7, check)
    m.close()
    assert n == on and np.array_equal(matches, omatches)
    assert n > 150 and (matches >= 0).sum() == n


def _contested_node_case(n_a, n_b, step, seed):
    """One vocabulary node in which n_a identical side-A descriptors compete for side-B descriptors at
    Hamming distances 0, step, 2 * step, ..: row i finds the i nearest candidates already taken, so the
    rows past the device's top-K list must be answered by the full scan of the row."""
    from orb_slam3_mac_b200._lib import KP_DTYPE
    from orb_slam3_mac_b200.matcher import BowSide
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 256, 32, dtype=np.uint8)
    da = np.repeat(x[None], n_a, axis=0)
    bits = np.unpackbits(np.repeat(x[None], n_b, axis=0), axis=1)
    for j in range(n_b):
        flip = min(j * step, 256)
        bits[j, :flip] ^= 1
    db = np.packbits(bits, axis=1)
    order = rng.permutation(n_b)            # candidates are not sorted by distance inside the node
    db = db[order]

    def keys(n):
        k = np.zeros(n, KP_DTYPE)
        k["x"], k["y"] = rng.uniform(20, 700, n), rng.uniform(20, 400, n)
        k["angle"] = rng.uniform(0, 359, n).astype(np.float32)
        return k
    A = BowSide(keys(n_a), da, {7: list(range(n_a))})
    B = BowSide(keys(n_b), db, {7: list(rng.permutation(n_b))})
    return A, B


@pytest.mark.parametrize("n_a,n_b,step", [(12, 10, 1), (12, 10, 9), (40, 70, 3), (6, 3, 2)])
def test_search_by_bow_contested_node_full_scan(mo, n_a, n_b, step):
    """Both SearchByBoW overloads when earlier rows have taken a row's nearest candidates (more of them than the
    K keys k_bow_topk returns), with rows longer than a warp (70 candidates) and shorter than K (3)."""
    from orb_slam3_mac_b200.matcher import ORBmatcher
    A, B = _contested_node_case(n_a, n_b, step, 3 * n_a + step)
    ok_a, ok_b = np.ones(n_a, np.uint8), np.ones(n_b, np.uint8)
    ok_b[::7] = 0
    for check in (False, True):
        m = ORBmatcher(0.8, check)
        n, matches = m.SearchByBoW(A, ok_a, B, -1)
        on, omatches = mo.search_by_bow(A, ok_a, B, -1, 0.8, check)
        assert n == on and np.array_equal(matches, omatches)
        n2, m12 = m.SearchByBoWKF(A, ok_a, B, ok_b, -1, n_b - 2)
        on2, om12 = mo.search_by_bow_kf(A, ok_a, B, ok_b, -1, n_b - 2, 0.8, check)
        m.close()
        assert n2 == on2 and np.array_equal(m12, om12)
        if not check and step < 9:
            assert n >= min(4, n_b)   # at least as many accepted rows as the top-K list is long


@pytest.mark.parametrize("n_a,n_b,step,refuse", [(5, 30, 1, 12), (5, 30, 0, 20), (4, 70, 0, 0), (3, 6, 2, 5), (6, 40, 1, 40)])
def test_search_for_triangulation_refused_candidates(mo, n_a, n_b, step, refuse):
    """SearchForTriangulation when the pair test refuses a row's nearest candidates (more of them than the keys
    k_bow_topk returns: full scan of the row), and with equal distances, where the reference's
    'dist > bestDist -> continue' scan keeps the LAST accepted candidate of the node."""
    from orb_slam3_mac_b200.matcher import ORBmatcher
    A, B = _contested_node_case(n_a, n_b, step, 11 * n_a + refuse)
    da = np.unpackbits(A.desc, axis=1).astype(np.int32)
    db = np.unpackbits(B.desc, axis=1).astype(np.int32)
    dist = (da[:, None, :] != db[None, :, :]).sum(axis=2)
    order = np.argsort(dist, axis=1, kind="stable")
    table = np.ones((n_a, n_b), np.uint8)
    for i in range(n_a):
        table[i, order[i, :refuse]] = 0      # the nearest `refuse` candidates fail the epipolar test
        table[i, order[i, refuse::3]] = 0    # and every third of the rest
    has1, has2 = np.zeros(n_a, np.uint8), np.zeros(n_b, np.uint8)
    has2[1::5] = 1
    for check in (False, True):
        m = ORBmatcher(0.8, check)
        n, m12 = m.SearchForTriangulation(A, has1, B, has2, pair_test=lambda a, b: table[a, b])
        m.close()
        on, om12 = mo.search_for_triangulation(A, has1, None, B, has2, None, False, None, None, table, None, None, check)
        assert n == on and np.array_equal(m12, om12)
    if refuse < n_b and step <= 1:
        assert on > 0 or check


class _ProductBackend:
    """The CUDA matcher behind the oracle's method names (tests/golden/match_cases.py)."""

    def __init__(self):
        from orb_slam3_mac_b200.matcher import ORBmatcher
        self.M = ORBmatcher

    def _m(self, ratio, check):
        return self.M(ratio, check)

    def search_by_projection_points(self, F, mps, md, th, far, th_far, ratio):
        m = self._m(ratio, True)
        r = m.SearchByProjection(F, mps, md, th, far, th_far)
        m.close()
        return r

    def search_by_projection_frames(self, F, lo, ld, th, fwd, bwd, mbf, check):
        m = self._m(0.8, check)
        r = m.SearchByProjectionFrames(F, lo, ld, th, fwd, bwd, mbf)
        m.close()
        return r

    def search_for_initialization(self, F0, F1, prev, win, ratio, check):
        m = self._m(ratio, check)
        r = m.SearchForInitialization(F0, F1, prev, win)
        m.close()
        return r

    def search_by_bow(self, KF, ok, F, nleft, ratio, check):
        m = self._m(ratio, check)
        r = m.SearchByBoW(KF, ok, F, nleft)
        m.close()
        return r

    def search_by_projection_points_stereo(self, F, R, mps, mpr, md, th, far, th_far, ratio):
        m = self._m(ratio, True)
        r = m.SearchByProjectionStereo(F, R, mps, mpr, md, th, far, th_far)
        m.close()
        return r

    def knn2(self, q, t):
        m = self._m(0.8, True)
        r = m.knn2(q, t)
        m.close()
        return r


def test_cuda_matcher_and_input_stage_against_committed_goldens():
    """The CUDA path against tests/golden/match_golden.json -- needs no oracle library: matcher hashes
    frozen from the CPU oracle, input-stage hashes recorded from the real OpenCV (cv2 4.13)."""
    import json
    import os
    from orb_slam3_mac_b200.extractor import ORBextractor
    from tests.golden import match_cases as mc
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "match_golden.json")))
    assert mc.run_matcher_cases(_ProductBackend()) == gold["matcher"]
    bgr, mx, my = mc.input_stage_case()
    g = gold["input_stage"]
    ex = ORBextractor(1000, 1.2, 8, 20, 7, max_size=(480, 752))
    ex.set_input_format("bgr", 15)
    ex.set_rectify_maps(mx, my)
    ex(bgr, (0, 0))
    assert mc.sha(ex.pyramid_level(0, (480, 752))) == g["bgr_rect_gray"]
    ex.set_input_format("gray", 15)
    ex(bgr[:, :, 1].copy(), (0, 0))
    assert mc.sha(ex.pyramid_level(0, (480, 752))) == g["rect_gray_only"]
    ex.set_rectify_maps(None, None)
    ex.set_input_format("rgb", 15)
    ex(bgr, (0, 0))
    assert mc.sha(ex.pyramid_level(0, (480, 752))) == g["rgb_gray"]
    ex.close()


class _ProductN3:
    """The CUDA matcher behind the oracle's N3 method names (tests/golden/n3_cases.py)."""

    def __init__(self):
        from orb_slam3_mac_b200.matcher import ORBmatcher
        self.M = ORBmatcher

    def _run(self, ratio, check, f):
        m = self.M(ratio, check)
        try:
            return f(m)
        finally:
            m.close()

    def search_by_bow_kf(self, KF1, ok1, KF2, ok2, nu1, nu2, ratio, check):
        return self._run(ratio, check, lambda m: m.SearchByBoWKF(KF1, ok1, KF2, ok2, nu1, nu2))

    def search_by_projection_sim3(self, KF, pts, d, th, ratio, taken):
        return self._run(0.8, True, lambda m: m.SearchByProjectionSim3(KF, pts, d, th, ratio, taken))

    def search_by_projection_reloc(self, F, pts, d, th, od, check, taken):
        return self._run(0.8, check, lambda m: m.SearchByProjectionReloc(F, pts, d, th, od, taken))

    def fuse(self, KF, pts, d, th, inv, off, depth=None, bf=0.0):   # depth / bf: how the case derived pts["ur"]
        return self._run(0.8, True, lambda m: m.Fuse(KF, pts, d, th, inv, off))

    def search_by_sim3(self, KF1, KF2, p1, d1, p2, d2, th):
        return self._run(0.8, True, lambda m: m.SearchBySim3(KF1, KF2, p1, d1, p2, d2, th))

    def search_for_triangulation(self, KF1, has1, st1, KF2, has2, st2, only_stereo, ep, sf2, pair_ok, F12, ls2, check,
                                 overload=1):   # both overloads are one entry point of the C ABI
        test = None if pair_ok is None else (lambda a, b: pair_ok[a, b])   # the camera model as a callback
        return self._run(0.8, check, lambda m: m.SearchForTriangulation(KF1, has1, KF2, has2, st1, st2, only_stereo, ep,
                                                                        sf2, test, F12, ls2))

    def distinctive_descriptors(self, d, gp):
        return self._run(0.8, True, lambda m: m.ComputeDistinctiveDescriptors(d, gp))

    def bow_transform(self, voc, desc, levelsup):
        def f(m):
            h = m.load_vocabulary(voc)
            try:
                return m.transform(h, desc, levelsup)
            finally:
                m.free_vocabulary(h)
        return self._run(0.8, True, f)


def test_n3_searches_against_committed_goldens_and_oracle(oracle_libs):
    """SURVEY 8(f) N3 through the C ABI: SearchByBoW(KF,KF), SearchByProjection(pKF,Scw,..) / reloc, Fuse x2,
    SearchBySim3, SearchForTriangulation (callback, pinhole F12 and coarse forms), ComputeDistinctiveDescriptors, and the
    DBoW2 transform of N4 (BowVector doubles bit-identical).
    Bit-exact against tests/golden/n3_golden.json (frozen from the reference's own code, make_golden_n3.py) and, array
    by array, against the reference-held oracle (oracle/_ref/liborbmref.so) run here."""
    import json
    import os
    from tests.golden import n3_cases as nc
    got = nc.run_n3_cases(_ProductN3())
    exp = nc.run_n3_cases(oracle_libs.MatchOracle("ref" if oracle_libs.have_mref() else "port"))
    for case, r in exp.items():
        for k, v in r.items():
            g = got[case][k]
            assert (np.array_equal(g, v) if isinstance(v, np.ndarray) else g == v), (case, k)
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "n3_golden.json")))
    assert nc.digest(got) == gold
    assert got["sim3"]["n"] > 300 and got["fuse_chi2"]["n"] > 200 and got["tri_pinhole"]["n"] > 50


def test_distinctive_descriptors_edge_cases(matcher):
    """Empty batch, empty groups, single observations, identical descriptors (all medians 0 -> first row)."""
    assert len(matcher.ComputeDistinctiveDescriptors(np.zeros((0, 32), np.uint8), np.zeros(1, np.int32))) == 0
    rng = np.random.default_rng(4)
    one = rng.integers(0, 256, (1, 32), dtype=np.uint8)
    d = np.concatenate([one, np.repeat(one, 5, axis=0), rng.integers(0, 256, (2, 32), dtype=np.uint8)])
    gp = np.array([0, 0, 1, 6, 6, 8], np.int32)
    assert matcher.ComputeDistinctiveDescriptors(d, gp).tolist() == [-1, 0, 0, -1, 0]


def test_n3_empty_and_invalid_inputs(matcher):
    """Ragged / empty inputs of the N3 and N4 entry points: no points, no valid points, empty keyframes, and the
    argument errors (levels outside the scale table, foreign vocabulary ids) are reported, not executed."""
    from orb_slam3_mac_b200._lib import KP_DTYPE, OrbxError
    from orb_slam3_mac_b200._lib_match import PROJ_POINT_DTYPE
    from orb_slam3_mac_b200.matcher import BowSide, Frame, Vocabulary
    rng = np.random.default_rng(1)
    k = np.zeros(50, KP_DTYPE)
    k["x"], k["y"] = rng.uniform(20, 700, 50), rng.uniform(20, 400, 50)
    d = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    F = Frame(k, d, 752, 480, SF)
    none = np.zeros(0, PROJ_POINT_DTYPE)
    assert matcher.SearchByProjectionSim3(F, none, np.zeros((0, 32), np.uint8), 5)[0] == 0
    assert matcher.Fuse(F, none, np.zeros((0, 32), np.uint8))[0] == 0
    pts = np.zeros(7, PROJ_POINT_DTYPE)            # all invalid: nothing may match
    n, match = matcher.SearchByProjectionReloc(F, pts, d[:7], 10.0, 100)
    assert n == 0 and (match == -1).all()
    E = Frame(np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8), 752, 480, SF)
    pts["valid"], pts["u"], pts["v"] = 1, 100, 100
    n, bi, bd = matcher.Fuse(E, pts, d[:7])        # empty keyframe
    assert n == 0 and (bi == -1).all() and (bd == 256).all()
    assert matcher.SearchBySim3(E, E, none, np.zeros((0, 32), np.uint8), none, np.zeros((0, 32), np.uint8), 7.5)[0] == 0
    pts["level"] = 8                                # outside the 8-entry scale table
    with pytest.raises(OrbxError):
        matcher.SearchByProjectionSim3(F, pts, d[:7], 5)
    A = BowSide(k, d, {})                           # empty FeatureVector on one side
    B = BowSide(k, d, {3: list(range(50))})
    assert matcher.SearchByBoWKF(A, np.ones(50, np.uint8), B, np.ones(50, np.uint8))[0] == 0
    assert matcher.SearchForTriangulation(B, np.ones(50, np.uint8), B, np.zeros(50, np.uint8))[0] == 0   # all tracked
    n, m12 = matcher.SearchForTriangulation(B, np.zeros(50, np.uint8), B, np.zeros(50, np.uint8))         # coarse, self
    assert n == 50 and (m12 == np.arange(50)).all()
    bad = Vocabulary([0, 2, 2, 2], [1, 9], np.zeros((3, 32), np.uint8), [0, 1, 1], [0, 0, 1], 1)   # child id 9 of 3
    with pytest.raises(OrbxError):
        matcher.load_vocabulary(bad)
    voc = Vocabulary([0, 2, 2, 2], [1, 2], d[:3], [0, 1.5, 2.5], [0, 0, 1], 1)
    h = matcher.load_vocabulary(voc)
    bi, bv, fn, fp, fi = matcher.transform(h, np.zeros((0, 32), np.uint8), 0)
    assert len(bi) == 0 and len(fn) == 0 and fp.tolist() == [0]
    bi, bv, fn, fp, fi = matcher.transform(h, d[1:3], 0)   # the two words themselves, levelsup 0 -> node = word node
    assert bi.tolist() == [0, 1] and np.allclose(bv, [1.5 / 4.0, 2.5 / 4.0]) and fn.tolist() == [1, 2]
    matcher.free_vocabulary(h)


def test_resident_frames_match_plain_frames(matcher, mo):
    """orbm_frame_upload: a frame kept on the device (keypoints, descriptors, grid built once) gives exactly the
    results of the per-call upload path and of the reference-held oracle, search after search."""
    from orb_slam3_mac_b200.matcher import ResidentFrame
    from tests.golden import match_cases as mc
    (ka, da), (kb, db) = mc.frames(7, n=1700)
    rng = np.random.default_rng(12)
    ur = np.where(rng.random(len(kb)) < 0.5, kb["x"] - rng.uniform(1, 40, len(kb)), -1.0).astype(np.float32)
    F = _frame(kb, db, 752, 480, u_right=ur)
    RF = ResidentFrame(matcher, F)
    RF.set_u_right(ur)
    try:
        mps, md = S.map_points_from_frame(ka, da, 752, 480, 6000, 5)
        lo, ld = S.last_frame_obs(ka, da, 5)
        obs0 = np.where(rng.random(len(kb)) < 0.2, rng.integers(0, 3, len(kb)), -1).astype(np.int32)
        for rep in range(2):   # the resident grid serves every search of the frame
            got = matcher.SearchByProjection(RF, mps, md, 3.0, True, 30.0, kp_obs=obs0)
            want = mo.search_by_projection_points(F, mps, md, 3.0, True, 30.0, 0.8, kp_obs=obs0)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
            got = matcher.SearchByProjectionFrames(RF, lo, ld, 7.0, False, False, 40.0, kp_obs=obs0)
            want = mo.search_by_projection_frames(F, lo, ld, 7.0, False, False, 40.0, True, kp_obs=obs0)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
        from orb_slam3_mac_b200._lib_match import AREA_QUERY_DTYPE
        q = np.zeros(500, AREA_QUERY_DTYPE)
        q["x"], q["y"], q["r"] = rng.uniform(0, 752, 500), rng.uniform(0, 480, 500), 30.0
        q["min_level"], q["max_level"], q["valid"] = -1, -1, 1
        a, b = matcher.area_candidates(RF, q, None, cap=500000), matcher.area_candidates(F, q, None, cap=500000)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    finally:
        RF.close()


def test_descriptor_banks_match_host_descriptors(matcher, mo):
    """orbm_bank: the map points' descriptors kept on the device and named by row (pMP->GetDescriptor() at
    src/ORBmatcher.cc:80 / :2031 becomes a row index).  Both bank searches, on a plain and on a resident frame, with
    rows in a shuffled bank, row i (rows = NULL) and a frame's own bank, equal the reference-held oracle fed the
    descriptors themselves; bad rows and foreign banks are refused."""
    from orb_slam3_mac_b200._lib import OrbxError
    from orb_slam3_mac_b200.matcher import DescriptorBank, ORBmatcher, ResidentFrame
    from tests.golden import match_cases as mc
    (ka, da), (kb, db) = mc.frames(11, n=1600)
    rng = np.random.default_rng(21)
    ur = np.where(rng.random(len(kb)) < 0.5, kb["x"] - rng.uniform(1, 40, len(kb)), -1.0).astype(np.float32)
    F = _frame(kb, db, 752, 480, u_right=ur)
    RF = ResidentFrame(matcher, F)
    RF.set_u_right(ur)
    mps, md = S.map_points_from_frame(ka, da, 752, 480, 5000, 4)
    lo, ld = S.last_frame_obs(ka, da, 4)
    obs0 = np.where(rng.random(len(kb)) < 0.2, rng.integers(0, 3, len(kb)), -1).astype(np.int32)
    # one bank for every map point of the "map": the search's points sit at shuffled rows of it
    cap = len(md) + len(ld) + 100
    bank = DescriptorBank(matcher, cap)
    rows_p = rng.permutation(cap)[:len(md)].astype(np.int32)
    left = np.setdiff1d(np.arange(cap), rows_p)
    rows_l = rng.permutation(left)[:len(ld)].astype(np.int32)
    try:
        assert bank.rows == cap
        all_rows = np.zeros((cap, 32), np.uint8)
        all_rows[rows_p] = md
        all_rows[rows_l] = ld
        bank.write(0, all_rows[:1000])      # written in pieces, as map points come and change
        bank.write(1000, all_rows[1000:])
        want_p = mo.search_by_projection_points(F, mps, md, 3.0, True, 30.0, 0.8, kp_obs=obs0)
        want_f = mo.search_by_projection_frames(F, lo, ld, 7.0, False, False, 40.0, True, kp_obs=obs0)
        assert want_p[0] > 100 and want_f[0] > 100
        for frame in (F, RF, RF):
            got = matcher.SearchByProjection(frame, mps, bank, 3.0, True, 30.0, kp_obs=obs0, rows=rows_p)
            assert got[0] == want_p[0] and np.array_equal(got[1], want_p[1]) and np.array_equal(got[2], want_p[2])
            got = matcher.SearchByProjectionFrames(frame, lo, bank, 7.0, False, False, 40.0, kp_obs=obs0, rows=rows_l)
            assert got[0] == want_f[0] and np.array_equal(got[1], want_f[1]) and np.array_equal(got[2], want_f[2])
        # a map point's descriptor changes (ComputeDistinctiveDescriptors): one row rewritten, the next search sees it
        md2 = md.copy()
        md2[7] = ~md2[7]
        bank.write(int(rows_p[7]), md2[7:8])
        got = matcher.SearchByProjection(RF, mps, bank, 3.0, True, 30.0, kp_obs=obs0, rows=rows_p)
        want2 = mo.search_by_projection_points(F, mps, md2, 3.0, True, 30.0, 0.8, kp_obs=obs0)
        assert got[0] == want2[0] and np.array_equal(got[1], want2[1])
        # rows = NULL: row i
        b2 = DescriptorBank(matcher, len(md))
        b2.write(0, md)
        got = matcher.SearchByProjection(RF, mps, b2, 1.0, False, 30.0)
        want = mo.search_by_projection_points(F, mps, md, 1.0, False, 30.0, 0.8)
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
        b2.close()
        # a resident frame's own descriptors as a bank: map points created from its keypoints (src/MapPoint.cc:109)
        RA = ResidentFrame(matcher, _frame(ka, da, 752, 480))
        sel = rng.permutation(len(ka))[:900].astype(np.int32)
        mps3, _ = S.map_points_from_frame(ka[sel], da[sel], 752, 480, 900, 6)
        # (map_points_from_frame permutes and flips bits: the test only needs SOME points whose descriptors are rows)
        got = matcher.SearchByProjection(RF, mps3, RA.bank(), 3.0, False, 30.0, rows=sel)
        want = mo.search_by_projection_points(F, mps3, da[sel], 3.0, False, 30.0, 0.8)
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
        # refusals
        bad = rows_p.copy()
        bad[3] = cap
        with pytest.raises(OrbxError):
            matcher.SearchByProjection(RF, mps, bank, 3.0, True, 30.0, rows=bad)
        bad[3] = -1
        with pytest.raises(OrbxError):
            matcher.SearchByProjectionFrames(RF, lo, bank, 7.0, rows=bad[:len(lo)])
        with pytest.raises(OrbxError):
            bank.write(cap - 1, md[:2])
        other = ORBmatcher(0.8, True, 0)
        try:
            with pytest.raises(OrbxError):
                other.SearchByProjection(F, mps, bank, 3.0, True, 30.0, rows=rows_p)
            orphan = DescriptorBank(other, 64)   # never closed: orbm_destroy frees what its context still owns
            orphan.write(0, md[:64])
        finally:
            other.close()
        orphan.close()                           # (a no-op once the context is gone)
        # the arena is clean after the refusals: the next search is right
        got = matcher.SearchByProjection(RF, mps, bank, 3.0, True, 30.0, kp_obs=obs0, rows=rows_p)
        assert got[0] == want2[0] and np.array_equal(got[1], want2[1])
        got = matcher.SearchByProjection(RF, mps, md2, 3.0, True, 30.0, kp_obs=obs0)
        assert got[0] == want2[0] and np.array_equal(got[1], want2[1])
        RA.close()
    finally:
        bank.close()
        RF.close()


def test_frame_from_device_undistorts_like_opencv(matcher, mo, oracle_libs):
    """orbm_frame_from_device: extractor results that never leave the device -> Frame::UndistortKeyPoints +
    ComputeImageBounds (src/Frame.cc:738-800) on the device -> grid -> search.  mvKeysUn and the bounds are
    bit-identical to the real cv2.undistortPoints (what the reference calls), the searches to the reference-held
    oracle on a Frame built from the same undistorted keypoints; without distortion the keypoints pass through."""
    import torch
    cv2 = pytest.importorskip("cv2")
    from orb_slam3_mac_b200.matcher import Frame, ResidentFrame
    from tests.golden import match_cases as mc
    (ka, da), (kb, db) = mc.frames(9, n=1500)
    d_keys = torch.from_numpy(kb.view(np.uint8).reshape(-1).copy()).cuda()
    d_desc = torch.from_numpy(db.reshape(-1).copy()).cuda()
    torch.cuda.synchronize()
    cam = (458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0)
    RF = ResidentFrame(matcher, d_keys=d_keys.data_ptr(), d_desc=d_desc.data_ptr(), n=len(kb), camera=cam, width=752,
                       height=480, scale_factors=SF)
    try:
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        dist = np.array(cam[4:8], np.float32)
        pts = np.stack([kb["x"], kb["y"]], axis=1).reshape(-1, 1, 2).astype(np.float32)
        want = cv2.undistortPoints(pts, K, dist, None, K).reshape(-1, 2)
        assert np.array_equal(RF.keys["x"].view(np.uint32), want[:, 0].view(np.uint32))
        assert np.array_equal(RF.keys["y"].view(np.uint32), want[:, 1].view(np.uint32))
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(RF.keys[f], kb[f])
        c = cv2.undistortPoints(np.array([[0, 0], [752, 0], [0, 480], [752, 480]], np.float32).reshape(-1, 1, 2), K, dist,
                                None, K).reshape(-1, 2)
        bounds = (min(c[0, 0], c[2, 0]), min(c[0, 1], c[1, 1]), max(c[1, 0], c[3, 0]), max(c[2, 1], c[3, 1]))
        assert tuple(float(b) for b in RF.bounds) == tuple(float(b) for b in bounds)
        assert np.array_equal(RF.desc, db)
        # the same frame as a host snapshot for the oracle
        F = Frame(RF.keys.copy(), db, 752, 480, SF, bounds=bounds)
        assert float(F.grid_w_inv) == float(RF.grid_w_inv) and float(F.grid_h_inv) == float(RF.grid_h_inv)
        mps, md = S.map_points_from_frame(ka, da, 752, 480, 5000, 2)
        got = matcher.SearchByProjection(RF, mps, md, 3.0, False, 30.0)
        want = mo.search_by_projection_points(F, mps, md, 3.0, False, 30.0, 0.8)
        assert got[0] == want[0] and got[0] > 100 and np.array_equal(got[1], want[1])
    finally:
        RF.close()
    RF = ResidentFrame(matcher, d_keys=d_keys.data_ptr(), d_desc=d_desc.data_ptr(), n=len(kb), camera=None, width=752,
                       height=480, scale_factors=SF)
    try:
        assert np.array_equal(RF.keys.view(np.uint8), kb.view(np.uint8))
        assert tuple(float(b) for b in RF.bounds) == (0.0, 0.0, 752.0, 480.0)
    finally:
        RF.close()


def _matcher_sweep(n, seed):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        w, h = int(rng.integers(200, 1400)), int(rng.integers(160, 900))
        out.append((i, w, h, int(rng.integers(300, 2500)), int(rng.integers(200, 8000)),
                    float(rng.choice([1.0, 2.0, 3.0, 5.0])), float(rng.choice([3.0, 7.0, 15.0, 40.0])),
                    bool(rng.integers(0, 2)), int(rng.integers(0, 3)), float(rng.choice([0.0, 20.0, 47.9]))))
    return out


# (ORBX_SWEEP_N / ORBX_SWEEP_SEED widen the sweep for a one-off soak; 200 further cases passed when this was written)
@pytest.mark.parametrize("i,w,h,nkp,nmp,th,thf,far,motion,mbf",
                         _matcher_sweep(int(os.environ.get("ORBX_SWEEP_N", "8")), int(os.environ.get("ORBX_SWEEP_SEED", "77"))))
def test_random_tracking_search_sweep(matcher, mo, i, w, h, nkp, nmp, th, thf, far, motion, mbf):
    """Both tracking searches on random frame sizes, keypoint / map-point counts, window factors, stereo gates and motion
    directions (fixed seed), on a plain frame, a resident frame and with the descriptors in a bank: identical to the
    reference-held oracle."""
    _tracking_case(matcher, mo, i, w, h, nkp, nmp, th, thf, far, motion, mbf)


def test_window_search_that_outgrows_lists_and_arena(mo):
    """Found by the soak (seed 9, case 3): on a FRESH context a 40-px frame-to-frame window yields ~200 k pairs -- more
    than the first guess of the lists (the search reruns) and more than the pinned arena holds (it grows).  The arena
    and its device mirror, which the rerun still reads its queries from, must not be replaced between the two passes."""
    from orb_slam3_mac_b200.matcher import ORBmatcher
    m = ORBmatcher(0.8, True, 0)
    try:
        _tracking_case(m, mo, 3, 1196, 376, 2463, 2918, 5.0, 40.0, False, 2, 47.9)
    finally:
        m.close()


def _tracking_case(matcher, mo, i, w, h, nkp, nmp, th, thf, far, motion, mbf):
    from orb_slam3_mac_b200.matcher import DescriptorBank, ResidentFrame
    from tests.golden import match_cases as mc
    rng = np.random.default_rng(1000 + i)
    (ka, da), (kb, db) = mc.frames(50 + i, n=nkp, w=w, h=h)
    ur = np.where(rng.random(len(kb)) < 0.5, kb["x"] - rng.uniform(1, 40, len(kb)), -1.0).astype(np.float32) if mbf else None
    F = _frame(kb, db, w, h, u_right=ur)
    mps, md = S.map_points_from_frame(ka, da, w, h, nmp, i)
    lo, ld = S.last_frame_obs(ka, da, i)
    obs0 = np.where(rng.random(len(kb)) < 0.2, rng.integers(0, 3, len(kb)), -1).astype(np.int32)
    fwd, bwd = motion == 1, motion == 2
    want_p = mo.search_by_projection_points(F, mps, md, th, far, 30.0, 0.8, kp_obs=obs0)
    want_f = mo.search_by_projection_frames(F, lo, ld, thf, fwd, bwd, mbf, True, kp_obs=obs0)
    RF = ResidentFrame(matcher, F)
    if ur is not None:
        RF.set_u_right(ur)
    bank = DescriptorBank(matcher, len(md) + len(ld))
    try:
        bank.write(0, md)
        bank.write(len(md), ld)
        rows_l = np.arange(len(md), len(md) + len(ld), dtype=np.int32)
        for frame, dp, dl, rl in ((F, md, ld, None), (RF, md, ld, None), (RF, bank, bank, rows_l)):
            got = matcher.SearchByProjection(frame, mps, dp, th, far, 30.0, kp_obs=obs0)
            assert got[0] == want_p[0] and np.array_equal(got[1], want_p[1]) and np.array_equal(got[2], want_p[2])
            got = matcher.SearchByProjectionFrames(frame, lo, dl, thf, fwd, bwd, mbf, kp_obs=obs0, rows=rl)
            assert got[0] == want_f[0] and np.array_equal(got[1], want_f[1]) and np.array_equal(got[2], want_f[2])
    finally:
        bank.close()
        RF.close()

