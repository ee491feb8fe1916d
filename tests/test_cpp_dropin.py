"""The C++ drop-in class (orb-slam3-mac_b200/host/ORBextractor.h, same shape as the reference's
include/ORBextractor.h:43-109) driven exactly like Frame::ExtractORB, compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp):
    exe = os.path.join(tmp, "dropin")
    pkg = os.path.join(ROOT, "orb-slam3-mac_b200")
    cmd = ["g++", "-std=c++11", "-O2", "-I" + os.path.join(pkg, "host"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "oracle", "cvshim"), os.path.join(ROOT, "tests", "cpp", "dropin_main.cpp"),
           "-L" + pkg, "-lorbx_b200", "-Wl,-rpath," + pkg, "-o", exe]
    subprocess.run(cmd, check=True)
    return exe


def test_dropin_compiles_against_the_cv_facade(tmp_path):
    """CPU check: the header has the reference's class shape and links against the C ABI."""
    _build(str(tmp_path))


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,nf,lap", [(480, 752, 1000, (0, 1000)), (512, 512, 1500, (0, 511)), (720, 1280, 2000, (0, 1000))])
def test_dropin_matches_oracle(tmp_path, oracle_libs, synth, h, w, nf, lap):
    from oracle.pyoracle import KP_DTYPE
    exe = _build(str(tmp_path))
    img = synth.textured_noise(h, w, 17)
    raw, out = str(tmp_path / "in.raw"), str(tmp_path / "out.bin")
    img.tofile(raw)
    level = 3
    subprocess.run([exe, str(w), str(h), str(nf), str(lap[0]), str(lap[1]), str(level), raw, out], check=True)
    b = open(out, "rb").read()
    hdr = np.frombuffer(b, np.int32, 6)
    mono, n, empty, nlev, lw, lh = [int(x) for x in hdr]
    off = 24
    sf = np.frombuffer(b, np.float32, 8, off); off += 32
    is2 = np.frombuffer(b, np.float32, 8, off); off += 32
    k = np.frombuffer(b, KP_DTYPE, n, off); off += 28 * n
    d = np.frombuffer(b, np.uint8, 32 * n, off).reshape(n, 32); off += 32 * n
    pyr = np.frombuffer(b, np.uint8, (lw + 38) * (lh + 38), off).reshape(lh + 38, lw + 38)

    kind = "ref" if oracle_libs.have_ref() else "port"
    ora = oracle_libs.OracleExtractor(nf, 1.2, 8, 20, 7, kind=kind)
    mono_r, k_r, d_r = ora(img, lap)
    assert empty == -1 and nlev == 8
    assert (mono, n) == (mono_r, len(k_r))
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(k[f], k_r[f]), f
    assert np.abs(k["angle"] - k_r["angle"]).max() <= 1e-4 * 180 / np.pi
    assert np.unpackbits(d ^ d_r).sum() <= 1e-3 * d.size * 8
    t = ora.tables()
    assert np.array_equal(sf, t["scale"]) and np.array_equal(is2, t["inv_sigma2"])
    assert np.array_equal(pyr, ora.level(level, 19))  # mvImagePyramid incl. the REFLECT_101 border


# ---- ORBmatcher: the product's C++ class against the reference's, on the same Frame / MapPoint objects ---------------
@pytest.fixture(scope="module")
def ref_and_dropin(oracle_libs):
    """oracle/_ref/liborbmref.so and liborbmdrop.so are ONE wrapper (oracle/ref_match_capi.cpp) and one set of reference
    sources (Frame.cc, MapPoint.cc, DBoW2) around two different `ORB_SLAM3::ORBmatcher` classes: the reference's
    src/ORBmatcher.cc, and orb-slam3-mac_b200/host/ORBmatcher.{h,cc} on top of the C ABI."""
    if not (oracle_libs.have_mref() and oracle_libs.have_mdrop()):
        pytest.skip("oracle/_ref/liborbmref.so / liborbmdrop.so need /root/reference to build")
    return oracle_libs.MatchOracle("ref"), oracle_libs.MatchOracle("drop")


def _same(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if isinstance(x, np.ndarray):
            assert np.array_equal(x, y)
        else:
            assert x == y


def test_matcher_dropin_library_is_built(oracle_libs):
    """CPU check: the drop-in class compiles against the reference's own headers (through the facade) and links."""
    if not os.path.isdir("/root/reference"):
        pytest.skip("needs /root/reference to build")
    oracle_libs.build("mdrop")
    assert oracle_libs.have_mdrop()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [2, 5])
def test_matcher_dropin_equals_reference_class(ref_and_dropin, synth, seed):
    """ORBmatcher::SearchByProjection (both tracking overloads, mono / rectified stereo / fisheye stereo, forward and
    backward motion, keypoints that already hold a map point), SearchForInitialization and DescriptorDistance: the
    reference's Frame objects end in the same state whichever class ran."""
    from orb_slam3_mac_b200 import synth as S
    from orb_slam3_mac_b200.matcher import Frame, RightHalf
    from tests.golden import match_cases as mc
    ref, drop = ref_and_dropin
    SF = mc.SF
    rng = np.random.default_rng(300 + seed)
    (ka, da), (kb, db) = mc.frames(seed, n=1500)
    ur = np.where(rng.random(len(kb)) < 0.6, kb["x"] - rng.uniform(1, 40, len(kb)), -1.0).astype(np.float32)
    FB, FBs = Frame(kb, db, 752, 480, SF), Frame(kb, db, 752, 480, SF, u_right=ur)
    obs0 = np.where(rng.random(len(kb)) < 0.2, rng.integers(0, 3, len(kb)), -1).astype(np.int32)
    for a, b in ((da[0], db[0]), (da[1], da[1]), (da[2], ~da[2])):
        assert ref.distance(a, b) == drop.distance(a, b)
    mps, md = S.map_points_from_frame(ka, da, 752, 480, 4000, seed)
    for F, th, far in ((FB, 1.0, False), (FBs, 3.0, True), (FBs, 1.0, False)):
        want = ref.search_by_projection_points(F, mps, md, th, far, 30.0, 0.8, kp_obs=obs0)
        _same(want, drop.search_by_projection_points(F, mps, md, th, far, 30.0, 0.8, kp_obs=obs0))
        assert want[0] > 100
    lo, ld = S.last_frame_obs(ka, da, seed)
    for F, th, fwd, bwd, mbf, check in ((FB, 15.0, False, False, 0.0, True), (FBs, 7.0, True, False, 40.0, True),
                                        (FBs, 7.0, False, True, 40.0, False)):
        want = ref.search_by_projection_frames(F, lo, ld, th, fwd, bwd, mbf, check, kp_obs=obs0)
        _same(want, drop.search_by_projection_frames(F, lo, ld, th, fwd, bwd, mbf, check, kp_obs=obs0))
        assert want[0] > 100
    k0, k1 = ka.copy(), kb.copy()
    k0["octave"] = np.where(rng.random(len(k0)) < 0.7, 0, 2)
    k1["octave"] = np.where(rng.random(len(k1)) < 0.8, 0, 1)
    F0, F1 = Frame(k0, da, 752, 480, SF), Frame(k1, db, 752, 480, SF)
    prev = np.stack([k0["x"] + 3, k0["y"]], axis=1).astype(np.float32)
    for window, ratio, check in ((10, 0.9, True), (100, 0.9, True), (40, 0.6, False)):
        _same(ref.search_for_initialization(F0, F1, prev, window, ratio, check),
              drop.search_for_initialization(F0, F1, prev, window, ratio, check))
    # the fisheye stereo forms (Nleft != -1)
    FA = Frame(ka, da, 752, 480, SF)
    R = RightHalf(kb, db, np.where(np.arange(len(ka)) % 3 == 0, np.arange(len(ka)) % len(kb), -1).astype(np.int32),
                  np.where(np.arange(len(kb)) % 4 == 0, np.arange(len(kb)) % len(ka), -1).astype(np.int32))
    uvr = np.stack([lo["u"] - 3, lo["v"]], axis=1).astype(np.float32)
    obs2 = np.where(rng.random(len(ka) + len(kb)) < 0.2, rng.integers(0, 3, len(ka) + len(kb)), -1).astype(np.int32)
    lo2 = lo.copy()
    lo2["u"] -= 3.0
    for th, fwd, bwd, check in ((15.0, False, False, True), (7.0, True, False, False), (7.0, False, True, True)):
        _same(ref.search_by_projection_frames_stereo(FA, R, lo2, uvr, ld, th, fwd, bwd, check, kp_obs=obs2),
              drop.search_by_projection_frames_stereo(FA, R, lo2, uvr, ld, th, fwd, bwd, check, kp_obs=obs2))
    from orb_slam3_mac_b200._lib_match import MAP_POINT_RIGHT_DTYPE
    m = 4000
    mpl, mdl = S.map_points_from_frame(ka, da, 752, 480, m, seed + 1)
    mr, mdr = S.map_points_from_frame(kb, db, 752, 480, m, seed + 2)
    use_r = rng.random(m) < 0.5        # every other map point carries the right camera's descriptor
    mdl = np.where(use_r[:, None], mdr, mdl)
    mpr = np.zeros(m, MAP_POINT_RIGHT_DTYPE)
    mpr["proj_x"], mpr["proj_y"], mpr["view_cos"] = mr["proj_x"], mr["proj_y"], mr["view_cos"]
    mpr["level"] = np.where(rng.random(m) < 0.05, -1, mr["level"])
    mpr["in_view"] = (rng.random(m) < 0.7).astype(np.uint8)
    mpl["in_view"] = (rng.random(m) < 0.7).astype(np.uint8)
    for th, far in ((1.0, False), (3.0, True)):
        want = ref.search_by_projection_points_stereo(FA, R, mpl, mpr, mdl, th, far, 30.0, 0.8, kp_obs=obs2)
        _same(want, drop.search_by_projection_points_stereo(FA, R, mpl, mpr, mdl, th, far, 30.0, 0.8, kp_obs=obs2))
        assert (want[1][: FA.N] >= 0).sum() > 50 and (want[1][FA.N:] >= 0).sum() > 50
    # SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches): mono and fisheye frame forms, node lists that partly overlap
    from orb_slam3_mac_b200.matcher import BowSide
    from tests.test_gpu_matcher import _feat_vec
    KFs = BowSide(ka, da, _feat_vec(ka, da, rng, drop=0.05))
    Fs = BowSide(kb, db, _feat_vec(kb, db, rng, drop=0.05))
    ok = (rng.random(len(ka)) < 0.7).astype(np.uint8)
    for nleft, ratio, check in ((-1, 0.7, True), (len(kb) // 2, 0.7, True), (-1, 0.9, False)):
        want = ref.search_by_bow(KFs, ok, Fs, nleft, ratio, check)
        _same(want, drop.search_by_bow(KFs, ok, Fs, nleft, ratio, check))
        assert want[0] > 50
    # SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12), incl. the fisheye keyframes' index limit (:863-865)
    ok2 = (rng.random(len(kb)) < 0.7).astype(np.uint8)
    for n_un1, n_un2, ratio, check in ((-1, -1, 0.8, True), (len(ka) - 200, len(kb) - 100, 0.8, True), (-1, -1, 0.9, False)):
        want = ref.search_by_bow_kf(KFs, ok, Fs, ok2, n_un1, n_un2, ratio, check)
        _same(want, drop.search_by_bow_kf(KFs, ok, Fs, ok2, n_un1, n_un2, ratio, check))
        assert want[0] > 50

