"""Seeded matcher / input-stage cases shared by the golden generator (make_golden_match.py), the CPU
test of the oracle and the GPU test of the CUDA path.  Every case returns plain arrays; the golden
file holds their sha1 (plus a few counts that make a mismatch readable)."""
import hashlib

import numpy as np

SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def frames(seed, n=1800, w=752, h=480):
    """Two related frames: B re-observes A's keypoints (3 px shift + noise, 3 % of the descriptor bits
    flipped) plus unrelated ones."""
    from orb_slam3_mac_b200 import synth as S
    from orb_slam3_mac_b200._lib import KP_DTYPE
    rng = np.random.default_rng(seed)
    ka = np.zeros(n, KP_DTYPE)
    ka["x"] = np.round(rng.uniform(20, w - 20, n))
    ka["y"] = np.round(rng.uniform(20, h - 20, n))
    ka["octave"] = rng.integers(0, 8, n)
    ka["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    da = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    kb = ka.copy()
    kb["x"] = np.round(ka["x"] + 3 + rng.normal(0, 1.0, n))
    kb["y"] = np.round(ka["y"] + rng.normal(0, 1.0, n))
    kb["angle"] = np.mod(ka["angle"] + rng.normal(0, 2.0, n), 360).astype(np.float32)
    db = S.flip_bits(da, 0.03, rng)
    fresh = rng.random(n) < 0.3
    db[fresh] = rng.integers(0, 256, (int(fresh.sum()), 32), dtype=np.uint8)
    perm = rng.permutation(n)
    return (ka, da), (kb[perm].copy(), db[perm].copy())


def feat_vec(keys, rng):
    node = np.floor(keys["x"] / 64).astype(np.int64) * 16 + np.floor(keys["y"] / 64).astype(np.int64)
    stray = rng.random(len(keys)) < 0.1
    node[stray] = rng.integers(0, 300, int(stray.sum()))
    fv = {}
    for i in rng.permutation(len(keys)):
        fv.setdefault(int(node[i]) * 5 + 1, []).append(int(i))
    return fv


def run_matcher_cases(backend):
    """backend: object with the oracle's method names (MatchOracle, or the adapter in the GPU test)."""
    from orb_slam3_mac_b200 import synth as S
    from orb_slam3_mac_b200._lib_match import MAP_POINT_RIGHT_DTYPE
    from orb_slam3_mac_b200.matcher import BowSide, Frame, RightHalf
    out = {}
    (ka, da), (kb, db) = frames(1)
    FA, FB = Frame(ka, da, 752, 480, SF), Frame(kb, db, 752, 480, SF)
    # M2
    mps, md = S.map_points_from_frame(ka, da, 752, 480, 5000, 3)
    n, match, obs = backend.search_by_projection_points(FB, mps, md, 3.0, True, 30.0, 0.8)
    out["points"] = {"n": int(n), "match": sha(match), "obs": sha(obs)}
    # M3
    lo, ld = S.last_frame_obs(ka, da, 4)
    lo["u"] -= 3.0
    n, match, obs = backend.search_by_projection_frames(FB, lo, ld, 9.0, False, False, 0.0, True)
    out["frames"] = {"n": int(n), "match": sha(match), "obs": sha(obs)}
    # M4
    k0, k1 = ka.copy(), kb.copy()
    k0["octave"] = np.where(np.arange(len(k0)) % 3 == 0, 1, 0)
    k1["octave"] = 0
    F0, F1 = Frame(k0, da, 752, 480, SF), Frame(k1, db, 752, 480, SF)
    prev = np.stack([k0["x"], k0["y"]], axis=1).astype(np.float32)
    n, m12, prev2 = backend.search_for_initialization(F0, F1, prev, 40, 0.9, True)
    out["init"] = {"n": int(n), "m12": sha(m12), "prev": sha(prev2)}
    # N3 SearchByBoW
    rng = np.random.default_rng(5)
    KF, FF = BowSide(ka, da, feat_vec(ka, rng)), BowSide(kb, db, feat_vec(kb, rng))
    ok = (rng.random(len(ka)) < 0.8).astype(np.uint8)
    n, m = backend.search_by_bow(KF, ok, FF, -1, 0.8, True)
    out["bow"] = {"n": int(n), "matches": sha(m)}
    n, m = backend.search_by_bow(KF, ok, FF, len(kb) // 2, 0.8, False)
    out["bow_fisheye"] = {"n": int(n), "matches": sha(m)}
    # M2 fisheye stereo: A's keypoints are the left camera, B's the right one
    R = RightHalf(kb, db, np.where(np.arange(len(ka)) % 4 == 0, np.arange(len(ka)) % len(kb), -1).astype(np.int32),
                  np.where(np.arange(len(kb)) % 5 == 0, np.arange(len(kb)) % len(ka), -1).astype(np.int32))
    mpl, mdl = S.map_points_from_frame(ka, da, 752, 480, 4000, 6)
    mr, mdr = S.map_points_from_frame(kb, db, 752, 480, 4000, 7)
    use_r = np.arange(4000) % 2 == 1
    mdl = np.where(use_r[:, None], mdr, mdl)
    mpr = np.zeros(4000, MAP_POINT_RIGHT_DTYPE)
    mpr["proj_x"], mpr["proj_y"], mpr["view_cos"], mpr["level"] = mr["proj_x"], mr["proj_y"], mr["view_cos"], mr["level"]
    mpr["in_view"] = (np.arange(4000) % 3 != 0).astype(np.uint8)
    n, match, obs = backend.search_by_projection_points_stereo(FA, R, mpl, mpr, mdl, 1.0, False, 30.0, 0.8)
    out["points_stereo"] = {"n": int(n), "match": sha(match), "obs": sha(obs)}
    # M6
    idx, dist = backend.knn2(da[:700], db[:900])
    out["knn2"] = {"idx": sha(idx), "dist": sha(dist)}
    return out


def input_stage_case():
    """A BGR frame and EuRoC-like rectification maps; the expected level 0 comes from the REAL OpenCV
    (cv2.remap then cv2.cvtColor) in make_golden_match.py."""
    from orb_slam3_mac_b200 import synth as S
    h, w = 480, 752
    rng = np.random.default_rng(77)
    base = S.frame("noise", h, w, 55)
    bgr = np.ascontiguousarray(np.stack(
        [np.clip(base.astype(np.int32) + rng.integers(-40, 41, (h, w)), 0, 255).astype(np.uint8) for _ in range(3)], axis=2))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    cx, cy = np.float32(367.2), np.float32(248.4)
    r2 = ((xx - cx) ** 2 + (yy - cy) ** 2) / np.float32(458.0 ** 2)
    k1 = np.float32(0.2)   # corners sample outside the image: BORDER_CONSTANT zeros
    mx = (cx + (xx - cx) * (1 + k1 * r2)).astype(np.float32)
    my = (cy + (yy - cy) * (1 + k1 * r2)).astype(np.float32)
    return bgr, mx, my
