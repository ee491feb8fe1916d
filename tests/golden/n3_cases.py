"""Seeded cases of the SURVEY 8(f) N3 searches (LocalMapping / LoopClosing / relocalisation), shared by
the golden generator (make_golden_n3.py), the CPU test of the oracle and the GPU test of the CUDA path.
run_n3_cases(backend) returns the raw result arrays of every case; golden file = their sha1 + counts."""
import numpy as np

from .match_cases import SF, feat_vec, frames, sha  # noqa: F401


BF = np.float32(40.0)   # the keyframes' mbf in the Fuse cases


def proj_points(keys, rng, jitter=1.5, invalid=0.15, level_shift=True):
    """Map points 'projected' next to the keypoints they came from (position + noise, predicted level =
    octave or octave + 1, so both ends of the [level-1, level] gate are used)."""
    from orb_slam3_mac_b200._lib_match import PROJ_POINT_DTYPE
    n = len(keys)
    p = np.zeros(n, PROJ_POINT_DTYPE)
    p["u"] = (keys["x"] + rng.normal(0, jitter, n)).astype(np.float32)
    p["v"] = (keys["y"] + rng.normal(0, jitter, n)).astype(np.float32)
    # the point's camera-frame depth; Fuse derives ur = u - bf * (1 / z) from it in float32 (src/ORBmatcher.cc:1448-1490)
    depth = rng.uniform(1.0, 20.0, n).astype(np.float32)
    p["ur"] = p["u"] - BF * (np.float32(1) / depth)
    p["angle"] = keys["angle"]
    lvl = keys["octave"].astype(np.int32)
    if level_shift:
        lvl = np.minimum(lvl + (rng.random(n) < 0.4), 7)
    p["level"] = lvl
    p["valid"] = (rng.random(n) >= invalid).astype(np.uint8)
    return p, depth


def n3_inputs(seed=11):
    from orb_slam3_mac_b200 import synth as S
    from orb_slam3_mac_b200.matcher import BowSide, Frame
    rng = np.random.default_rng(seed)
    (ka, da), (kb, db) = frames(seed, n=1600)
    # B's keypoints re-observe A's: a[perm-inverse]; recover the correspondence by nearest position
    ur_b = np.where(rng.random(len(kb)) < 0.5, kb["x"] - rng.uniform(2, 40, len(kb)), -1).astype(np.float32)
    FA = Frame(ka, da, 752, 480, SF)
    FB = Frame(kb, db, 752, 480, SF, u_right=ur_b)
    FBm = Frame(kb, db, 752, 480, SF)
    # points of A projected into B (B = A shifted by 3 px in x) and the reverse
    pa, za = proj_points(ka, rng)
    pa["u"] += 3.0
    pa["ur"] += 3.0
    pb, _ = proj_points(kb, rng)
    pb["u"] -= 3.0
    mpd_a = S.flip_bits(da, 0.04, rng)   # map-point descriptors differ a little from the keyframe's
    mpd_b = S.flip_bits(db, 0.04, rng)
    fva, fvb = feat_vec(ka, rng), feat_vec(kb, rng)
    KA, KB = BowSide(ka, da, fva), BowSide(kb, db, fvb)
    pa["ur"] = pa["u"] - BF * (np.float32(1) / za)   # after the shift: exactly what the reference computes from (u, bf, z)
    return dict(rng=rng, ka=ka, da=da, kb=kb, db=db, FA=FA, FB=FB, FBm=FBm, pa=pa, pb=pb, mpd_a=mpd_a, mpd_b=mpd_b,
                KA=KA, KB=KB, za=za)


def distinctive_case(seed=3, sizes=None):
    """Observation groups of 0..70 descriptors: noisy copies of one descriptor plus outliers (ties between
    medians are common, which exercises the first-minimum rule); groups on both sides of the 256-row limit
    of the register path."""
    from orb_slam3_mac_b200 import synth as S
    rng = np.random.default_rng(seed)
    if sizes is None:
        sizes = list(rng.integers(0, 25, 300)) + [1, 2, 3, 33, 64, 70, 150, 0, 256, 257, 300]
    rows, ptr = [], [0]
    for n in sizes:
        if n:
            base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
            g = S.flip_bits(np.repeat(base, n, axis=0), 0.06, rng)
            out = rng.random(n) < 0.15
            g[out] = rng.integers(0, 256, (int(out.sum()), 32), dtype=np.uint8)
            rows.append(g)
        ptr.append(ptr[-1] + int(n))
    return np.concatenate(rows), np.array(ptr, np.int32)


def run_n3_cases(backend):
    """backend: object with the oracle's N3 method names.  Returns {case: {name: array or int}}."""
    I = n3_inputs()
    rng = np.random.default_rng(99)
    ka, kb, FA, FB, FBm = I["ka"], I["kb"], I["FA"], I["FB"], I["FBm"]
    out = {}
    ok1 = (rng.random(len(ka)) < 0.8).astype(np.uint8)
    ok2 = (rng.random(len(kb)) < 0.8).astype(np.uint8)
    for name, check, nu1, nu2 in (("bow_kf", True, -1, -1), ("bow_kf_nocheck_nleft", False, 1200, 1300)):
        n, m12 = backend.search_by_bow_kf(I["KA"], ok1, I["KB"], ok2, nu1, nu2, 0.8, check)
        out[name] = {"n": int(n), "m12": m12}
    taken = (rng.random(len(kb)) < 0.2).astype(np.uint8)
    for name, th, ratio in (("sim3_proj", 8, 1.0), ("sim3_proj_tight", 4, 0.8)):
        n, match = backend.search_by_projection_sim3(FBm, I["pa"], I["mpd_a"], th, ratio, taken)
        out[name] = {"n": int(n), "match": match}
    for name, th, od, check in (("reloc", 10.0, 100, True), ("reloc_narrow", 3.0, 64, False)):
        n, match = backend.search_by_projection_reloc(FBm, I["pa"], I["mpd_a"], th, od, check, taken)
        out[name] = {"n": int(n), "match": match}
    inv_sigma2 = (1.0 / (SF * SF)).astype(np.float32)
    n, bi, bd = backend.fuse(FB, I["pa"], I["mpd_a"], 3.0, inv_sigma2, 0, depth=I["za"], bf=float(BF))
    out["fuse_chi2"] = {"n": int(n), "idx": bi, "dist": bd}
    n, bi, bd = backend.fuse(FBm, I["pa"], I["mpd_a"], 4.0, None, 5000)
    out["fuse_sim3"] = {"n": int(n), "idx": bi, "dist": bd}
    n, m12 = backend.search_by_sim3(FA, FBm, I["pa"], I["mpd_a"], I["pb"], I["mpd_b"], 7.5)
    out["sim3"] = {"n": int(n), "m12": m12}
    # triangulation: camera-model test as a table / pinhole F12 / coarse
    has1 = (rng.random(len(ka)) < 0.5).astype(np.uint8)
    has2 = (rng.random(len(kb)) < 0.5).astype(np.uint8)
    st1 = (rng.random(len(ka)) < 0.5).astype(np.uint8)
    st2 = (rng.random(len(kb)) < 0.5).astype(np.uint8)
    pair_ok = (rng.random((len(ka), len(kb))) < 0.7).astype(np.uint8)
    ep = np.array([400.0, 250.0], np.float32)
    ls2 = (SF * SF).astype(np.float32)
    # F12 of a pure x-translation between identical pinhole cameras: epipolar lines are image rows
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) * np.float32(0.37)
    n, m12 = backend.search_for_triangulation(I["KA"], has1, st1, I["KB"], has2, st2, False, ep, SF, pair_ok, None, None, True)
    out["tri_table"] = {"n": int(n), "m12": m12}
    n, m12 = backend.search_for_triangulation(I["KA"], has1, None, I["KB"], has2, None, False, ep, SF, None, F12, ls2, True)
    out["tri_pinhole"] = {"n": int(n), "m12": m12}
    n, m12 = backend.search_for_triangulation(I["KA"], has1, st1, I["KB"], has2, st2, True, None, None, None, None, None, False)
    out["tri_coarse_stereo"] = {"n": int(n), "m12": m12}
    # the matchAndtriangulate overload (src/ORBmatcher.cc:1212-1401): no stereo / epipole gates
    n, m12 = backend.search_for_triangulation(I["KA"], has1, None, I["KB"], has2, None, False, None, None, pair_ok, None, None, True,
                                              overload=2)
    out["tri_overload2"] = {"n": int(n), "m12": m12}
    d, gp = distinctive_case()
    out["distinctive"] = {"best": backend.distinctive_descriptors(d, gp)}
    # N4: DBoW2 transform on a synthetic vocabulary (ragged and regular trees, every weighting / norm branch)
    for name, ragged, L, levelsup, weighting, norm in (("bow_ragged_l2", True, 4, 2, 0, 1), ("bow_root", True, 4, 4, 0, 1),
                                                       ("bow_regular_tf", False, 3, 1, 1, 0),
                                                       ("bow_binary_l2", False, 3, 2, 3, 2), ("bow_idf", True, 4, 3, 2, 1)):
        voc = synthetic_vocabulary(21 + L, 10, L, ragged)
        voc.weighting, voc.norm = weighting, norm
        f = bow_features(voc, 1500)
        bi, bv, fn, fp, fi = backend.bow_transform(voc, f, levelsup)
        out[name] = {"n_words": int(len(bi)), "n_nodes": int(len(fn)), "bow_ids": bi, "bow_vals": bv, "fv_nodes": fn,
                     "fv_ptr": fp, "fv_items": fi}
    return out


def digest(res):
    return {c: {k: (sha(v) if isinstance(v, np.ndarray) else v) for k, v in r.items()} for c, r in res.items()}


def synthetic_vocabulary(seed=21, k=10, L=4, ragged=True):
    """A DBoW2-shaped vocabulary (ORBvoc.txt is not available offline): k-ary tree of depth L built breadth first as
    DBoW2 numbers its nodes, child descriptors = parent's with bits flipped (so descents are decided by small
    distance differences and ties occur), random idf weights with a few stopped words (weight 0).  ragged: some
    inner nodes have fewer than k children and some branches end early (words above level L)."""
    from orb_slam3_mac_b200 import synth as S
    from orb_slam3_mac_b200.matcher import Vocabulary
    rng = np.random.default_rng(seed)
    desc = [rng.integers(0, 256, 32, dtype=np.uint8)]
    children = [[]]
    level = [0]
    frontier = [0]
    for lv in range(1, L + 1):
        nxt = []
        for p in frontier:
            if ragged and lv > 1 and rng.random() < 0.05:
                continue   # p stays a leaf: a word above the last level
            nc = k if not ragged or rng.random() < 0.8 else int(rng.integers(1, k))
            for _ in range(nc):
                nid = len(desc)
                desc.append(S.flip_bits(desc[p][None, :], 0.12 if lv > 1 else 0.5, rng)[0])
                children.append([])
                level.append(lv)
                children[p].append(nid)
                nxt.append(nid)
        frontier = nxt
    n = len(desc)
    child_ptr = np.zeros(n + 1, np.int32)
    ids = []
    for i in range(n):
        ids.extend(children[i])
        child_ptr[i + 1] = len(ids)
    weight = np.zeros(n, np.float64)
    word_id = np.zeros(n, np.uint32)
    leaves = [i for i in range(n) if not children[i]]
    for w, i in enumerate(leaves):
        word_id[i] = w
        weight[i] = 0.0 if rng.random() < 0.03 else float(rng.uniform(0.5, 9.0))
    return Vocabulary(child_ptr, np.array(ids, np.uint32), np.stack(desc), weight, word_id, L)


def bow_features(voc, n, seed=5):
    """Descriptors near random vocabulary nodes (plus pure noise), duplicates included so words repeat."""
    from orb_slam3_mac_b200 import synth as S
    rng = np.random.default_rng(seed)
    pick = rng.integers(1, len(voc.node_desc), n)
    d = S.flip_bits(voc.node_desc[pick], 0.05, rng)
    noise = rng.random(n) < 0.1
    d[noise] = rng.integers(0, 256, (int(noise.sum()), 32), dtype=np.uint8)
    d[n // 2:n // 2 + 40] = d[:40]
    return d
