#!/usr/bin/env python
"""Measurement of the matcher rows of SURVEY.md section 8 (M1-M7, N1, N2, N3, N4's DBoW2 transform) -- companion of bench.py,
which measures the extractor and carries the driver's JSON contract.

For every case it prints one JSON line:
  pairs           descriptor pairs the call compares (8 popc each)
  device_ms       CUDA-event time around the call's kernels (orbm_last_device_ms), median
  e2e_ms          wall time of the C-ABI call with HOST buffers (H2D + kernels + D2H + the host accept
                  loop of the order-dependent searches), median
  cpu_ms          the CPU oracle (oracle/match_oracle.cpp, one thread, the reference's loops) on the
                  same snapshot, on this box's host cores
  gpairs_per_s    pairs / device_ms; popc_frac = 8 * that / the measured POPC issue rate
Inputs are frames extracted by the CUDA extractor from seeded synthetic images (synth.py).
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def med(f, reps):
    ts = []
    out = None
    for _ in range(reps):
        t0 = time.perf_counter()
        out = f()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e3, out


ROWS = []


def collect(reps: int = 15, cpu: bool = True):
    """The rows as a list (bench.py puts them under `matcher` in its JSON line)."""
    ROWS.clear()
    main(["--reps", str(reps), "--quiet"] + ([] if cpu else ["--no-cpu"]))
    return list(ROWS)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--quiet", action="store_true")
    args = ap.parse_args(argv)

    importlib.import_module("orb_slam3_mac_b200")
    from orb_slam3_mac_b200 import synth as S
    from orb_slam3_mac_b200.extractor import ORBextractor
    from orb_slam3_mac_b200.matcher import Frame, ORBmatcher
    from orb_slam3_mac_b200._lib_match import AREA_QUERY_DTYPE
    mo = None
    if not args.no_cpu:
        from oracle import pyoracle
        pyoracle.build("port")
        mo = pyoracle.MatchOracle()   # the timed CPU arm: our port (the verbatim build is the parity checker)

    m = ORBmatcher(0.8, True)
    lib, ctx = m.lib, m.ctx
    peak = C.c_double()
    lib.orbm_measure_popc_peak(ctx, C.byref(peak))
    SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)

    def dev_ms():
        ms, pairs = C.c_float(), C.c_longlong()
        lib.orbm_last_device_ms(ctx, C.byref(ms), C.byref(pairs))
        return float(ms.value), int(pairs.value)

    def report(case, what, call, cpu_call):
        call()  # warm-up (buffer growth)
        dms = []
        e2e, _ = med(lambda: (call(), dms.append(dev_ms()))[0], args.reps)
        d = float(np.median([x[0] for x in dms]))
        pairs = dms[-1][1]
        cpu = None
        if cpu_call is not None and mo is not None:
            cpu, _ = med(cpu_call, 3)
        gp = pairs / (d * 1e-3) * 1e-9 if d > 0 else None
        row = {"case": case, "what": what, "pairs": pairs, "device_ms": round(d, 4),
               "e2e_ms": round(e2e, 4), "cpu_ms": None if cpu is None else round(cpu, 3),
               "cpu_kind": "port, 1 thread", "gpairs_per_s": None if gp is None else round(gp, 2),
               "popc_peak_gpopc_per_s": round(peak.value, 1),
               "popc_frac": None if gp is None else round(8 * gp / peak.value, 4)}
        ROWS.append(row)
        if not args.quiet:
            print(json.dumps(row), flush=True)

    def extract(img, nf):
        ex = ORBextractor(nf, 1.2, 8, 20, 7, max_size=img.shape)
        k, d = ex(img, (0, 0))[1:3]
        return k, d

    # ---- M6: fisheye stereo brute force 2-NN at C3 (1500 x 1500) and a large synthetic case ------
    imgL, imgR = S.fisheye_pair(512, 5)
    kL, dL = extract(imgL, 1500)
    kR, dR = extract(imgR, 1500)
    report("C3 knn2", f"cv::BFMatcher knnMatch k=2, {len(dL)} x {len(dR)} descriptors", lambda: m.knn2(dL, dR),
           (lambda: mo.knn2(dL, dR)) if mo else None)
    rng = np.random.default_rng(0)
    bigq = rng.integers(0, 256, (16384, 32), dtype=np.uint8)
    bigt = rng.integers(0, 256, (16384, 32), dtype=np.uint8)
    report("knn2 16k x 16k", "same kernel at a size that fills the GPU", lambda: m.knn2(bigq, bigt), None)

    # ---- M5 + N1: rectified stereo at C2 -------------------------------------------------------------
    iL, iR = S.stereo_pair(480, 752, 3)
    exL, exR = ORBextractor(1200, 1.2, 8, 20, 7), ORBextractor(1200, 1.2, 8, 20, 7)
    kl, dl = exL(iL, (0, 0))[1:3]
    kr, dr = exR(iR, (0, 0))[1:3]
    FL, FR = Frame(kl, dl, 752, 480, SF), Frame(kr, dr, 752, 480, SF)
    report("C2 stereo rows", f"ComputeStereoMatches Hamming arg-min, {len(kl)} x {len(kr)}",
           lambda: m.stereo_rows(FL, FR, 0.0, 60.0), (lambda: mo.stereo_rows(FL, FR, 480, 0.0, 60.0)) if mo else None)
    pl, pr = exL.device_pyramid(), exR.device_pyramid()
    report("C2 ComputeStereoMatches", "rows + SAD refinement on the device-resident pyramids + median cut",
           lambda: m.ComputeStereoMatches(FL, FR, pl, pr, 0.11, 47.9), None)

    # ---- M2 + M7/N2: SearchByProjection(Frame, map points) at C4 ---------------------------------------
    img0 = S.textured_noise(720, 1280, 7)
    img1 = np.roll(img0, 3, axis=1)
    k0, d0 = extract(img0, 2000)
    k1, d1 = extract(img1, 2000)
    mps, md = S.map_points_from_frame(k0, d0, 1280, 720, 10000, 7)
    F = Frame(k1, d1, 1280, 720, SF)
    report("C4 SearchByProjection", f"{len(mps)} map points vs {len(k1)} keypoints, th=1 (device grid + windows + distances, host accept loop)",
           lambda: m.SearchByProjection(F, mps, md, 1.0, False, 30.0),
           (lambda: mo.search_by_projection_points(F, mps, md, 1.0, False, 30.0, 0.8)) if mo else None)
    report("C4 SearchByProjection th=3", "same, 3x window",
           lambda: m.SearchByProjection(F, mps, md, 3.0, False, 30.0),
           (lambda: mo.search_by_projection_points(F, mps, md, 3.0, False, 30.0, 0.8)) if mo else None)

    # the same searches on a device-resident frame (orbm_frame: keypoints, descriptors and grid uploaded / built once)
    from orb_slam3_mac_b200.matcher import ResidentFrame
    RF = ResidentFrame(m, F)
    report("C4 SearchByProjection, resident frame", "same search, the frame stays on the device (no frame upload, no grid build)",
           lambda: m.SearchByProjection(RF, mps, md, 1.0, False, 30.0), None)
    lo_r, ld_r = S.last_frame_obs(k0, d0, 9)
    report("C4 SearchByProjection frames, resident frame", "frame-to-frame search on the resident frame",
           lambda: m.SearchByProjectionFrames(RF, lo_r, ld_r, 15.0), None)
    # ... and the map points' descriptors in a device-resident bank (orbm_bank): a call sends its window queries only
    from orb_slam3_mac_b200.matcher import DescriptorBank
    bank = DescriptorBank(m, len(md) + len(ld_r))
    bank.write(0, md)
    bank.write(len(md), ld_r)
    rows_l = np.arange(len(md), len(md) + len(ld_r), dtype=np.int32)
    want = m.SearchByProjection(RF, mps, md, 1.0, False, 30.0)
    got = m.SearchByProjection(RF, mps, bank, 1.0, False, 30.0)
    assert got[0] == want[0] and np.array_equal(got[1], want[1]), "bank search differs"
    report("C4 SearchByProjection, resident frame + descriptor bank", "same search, map-point descriptors resident too",
           lambda: m.SearchByProjection(RF, mps, bank, 1.0, False, 30.0), None)
    report("C4 SearchByProjection frames, resident frame + descriptor bank", "frame-to-frame search, descriptors resident too",
           lambda: m.SearchByProjectionFrames(RF, lo_r, bank, 15.0, rows=rows_l), None)
    bank.close()
    RF.close()

    # ---- M4: SearchForInitialization with the 5 x nFeatures extractor at C1 -----------------------------
    a = S.textured_noise(480, 752, 21)
    b = np.roll(a, 4, axis=1)
    ka, da = extract(a, 5000)
    kb, db = extract(b, 5000)
    F1, F2 = Frame(ka, da, 752, 480, SF), Frame(kb, db, 752, 480, SF)
    prev = np.stack([ka["x"], ka["y"]], axis=1).astype(np.float32)
    report("C1 SearchForInitialization", f"{len(ka)} x {len(kb)} keypoints, 100-px windows at level 0",
           lambda: m.SearchForInitialization(F1, F2, prev.copy(), 100),
           (lambda: mo.search_for_initialization(F1, F2, prev.copy(), 100, 0.8, True)) if mo else None)

    # ---- M3: SearchByProjection(CurrentFrame, LastFrame) ----------------------------------------------
    lo, ld = S.last_frame_obs(k0, d0, 9)
    report("C4 SearchByProjection frames", f"{len(lo)} last-frame observations vs {len(k1)} keypoints, th=15",
           lambda: m.SearchByProjectionFrames(F, lo, ld, 15.0),
           (lambda: mo.search_by_projection_frames(F, lo, ld, 15.0, False, False, 0.0, True)) if mo else None)

    # ---- N3: SearchByBoW(KeyFrame*, Frame&) on the tracking thread ------------------------------------------
    from orb_slam3_mac_b200.matcher import BowSide

    def feat_vec(keys, rng):   # stand-in for DBoW2's FeatureVector: ~100 nodes, node = coarse image cell
        node = np.floor(keys["x"] / 128).astype(np.int64) * 10 + np.floor(keys["y"] / 72).astype(np.int64)
        fv = {}
        for i in rng.permutation(len(keys)):
            fv.setdefault(int(node[i]), []).append(int(i))
        return fv

    rb = np.random.default_rng(3)
    KFs, Fs = BowSide(k0, d0, feat_vec(k0, rb)), BowSide(k1, d1, feat_vec(k1, rb))
    okf = np.ones(len(k0), np.uint8)
    report("C4 SearchByBoW", f"{len(k0)} keyframe features vs {len(k1)} frame features over {len(KFs.node_ids)} vocabulary nodes",
           lambda: m.SearchByBoW(KFs, okf, Fs, -1),
           (lambda: mo.search_by_bow(KFs, okf, Fs, -1, 0.8, True)) if mo else None)

    # ---- N3: LocalMapping / LoopClosing / relocalisation searches (C4-sized keyframes) ------------------
    from tests.golden import n3_cases as nc
    rp = np.random.default_rng(12)
    KF1, KF2 = Frame(k0, d0, 1280, 720, SF), Frame(k1, d1, 1280, 720, SF)
    p1, p2 = nc.proj_points(k0, rp)[0], nc.proj_points(k1, rp)[0]
    md1, md2 = S.flip_bits(d0, 0.04, rp), S.flip_bits(d1, 0.04, rp)
    ok1, ok2 = np.ones(len(k0), np.uint8), np.ones(len(k1), np.uint8)
    report("C4 SearchByBoW(KF,KF)", f"{len(k0)} x {len(k1)} keyframe features over {len(KFs.node_ids)} vocabulary nodes",
           lambda: m.SearchByBoWKF(KFs, ok1, Fs, ok2),
           (lambda: mo.search_by_bow_kf(KFs, ok1, Fs, ok2, -1, -1, 0.8, True)) if mo else None)
    report("C4 SearchByProjection(pKF,Scw)", f"{len(p1)} projected map points vs {len(k1)} keypoints, th=8",
           lambda: m.SearchByProjectionSim3(KF2, p1, md1, 8, 1.0),
           (lambda: mo.search_by_projection_sim3(KF2, p1, md1, 8, 1.0)) if mo else None)
    report("C4 SearchByProjection reloc", f"{len(p1)} keyframe map points vs {len(k1)} keypoints, th=10, ORBdist=100",
           lambda: m.SearchByProjectionReloc(KF2, p1, md1, 10.0, 100),
           (lambda: mo.search_by_projection_reloc(KF2, p1, md1, 10.0, 100, True)) if mo else None)
    inv2 = (1.0 / (SF * SF)).astype(np.float32)
    report("C4 Fuse", f"{len(p1)} map points into a keyframe of {len(k1)} keypoints, th=3, chi-square gates",
           lambda: m.Fuse(KF2, p1, md1, 3.0, inv2),
           (lambda: mo.fuse(KF2, p1, md1, 3.0, inv2, 0)) if mo else None)
    report("C4 SearchBySim3", f"{len(p1)} + {len(p2)} projected map points, both directions, th=7.5",
           lambda: m.SearchBySim3(KF1, KF2, p1, md1, p2, md2, 7.5),
           (lambda: mo.search_by_sim3(KF1, KF2, p1, md1, p2, md2, 7.5)) if mo else None)
    F12 = (np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32) * np.float32(0.37))
    ls2 = (SF * SF).astype(np.float32)
    no1, no2 = np.zeros(len(k0), np.uint8), np.zeros(len(k1), np.uint8)
    ep = np.array([640.0, 360.0], np.float32)
    report("C4 SearchForTriangulation", f"{len(k0)} x {len(k1)} untracked features, pinhole epipolar test",
           lambda: m.SearchForTriangulation(KFs, no1, Fs, no2, None, None, False, ep, SF, None, F12, ls2),
           (lambda: mo.search_for_triangulation(KFs, no1, None, Fs, no2, None, False, ep, SF, None, F12, ls2, True)) if mo else None)
    rs = np.random.default_rng(6)   # observation counts of a mature map: mostly 2..25, a tail up to 150
    sizes = list(rs.integers(2, 26, 300)) + [40, 64, 100, 150]
    dd, gp = nc.distinctive_case(5, sizes)
    reps_g = 40   # ~12 k map points per call
    dd = np.concatenate([dd] * reps_g)
    gp = np.concatenate([[0]] + [gp[1:] + k * gp[-1] for k in range(reps_g)]).astype(np.int32)
    report("ComputeDistinctiveDescriptors", f"{len(gp) - 1} map points, {len(dd)} observed descriptors (2..150 per point)",
           lambda: m.ComputeDistinctiveDescriptors(dd, gp),
           (lambda: mo.distinctive_descriptors(dd, gp)) if mo else None)

    # ---- N4: DBoW2 transform with an ORBvoc-shaped vocabulary (k = 10, L = 6: 1.1 M nodes, 36 MB of descriptors) --
    from orb_slam3_mac_b200.matcher import Vocabulary
    rv = np.random.default_rng(31)
    K, LV = 10, 6
    descs = [rv.integers(0, 256, (1, 32), dtype=np.uint8)]
    for lv in range(1, LV + 1):   # breadth first, as DBoW2 numbers its nodes
        par = np.repeat(descs[-1], K, axis=0)
        flip = np.packbits(rv.random((len(par), 256)) < (0.5 if lv == 1 else 0.12), axis=1)
        descs.append(par ^ flip)
    nd = np.concatenate(descs)
    n_nodes = len(nd)
    first = np.cumsum([0] + [len(d) for d in descs])      # first node id of each level
    n_inner = int(first[LV])
    child_ptr = np.concatenate([np.arange(n_inner + 1) * K, np.full(n_nodes - n_inner, n_inner * K)]).astype(np.int32)
    child_ids = np.arange(1, n_nodes, dtype=np.uint32)     # node i's children are 1 + K i .. K + K i
    weight = np.zeros(n_nodes)
    weight[n_inner:] = rv.uniform(0.5, 9.0, n_nodes - n_inner)
    word_id = np.zeros(n_nodes, np.uint32)
    word_id[n_inner:] = np.arange(n_nodes - n_inner)
    voc = Vocabulary(child_ptr, child_ids, nd, weight, word_id, LV)
    vh = m.load_vocabulary(voc)
    report("C4 DBoW2 transform", f"{len(d1)} descriptors through a k={K}, L={LV} vocabulary ({n_nodes} nodes), levelsup 4",
           lambda: m.transform(vh, d1, 4), (lambda: mo.bow_transform(voc, d1, 4)) if mo else None)
    m.free_vocabulary(vh)

    # ---- N2 alone: candidate lists without distances ---------------------------------------------------
    q = np.zeros(len(mps), AREA_QUERY_DTYPE)
    q["x"], q["y"], q["r"] = mps["proj_x"], mps["proj_y"], 4.0 * SF[np.clip(mps["level"], 0, 7)]
    q["min_level"], q["max_level"], q["valid"] = mps["level"] - 1, mps["level"], 1
    report("C4 grid + windows", f"AssignFeaturesToGrid + {len(q)} GetFeaturesInArea queries on the device",
           lambda: m.area_candidates(F, q, None, cap=200_000), None)
    m.close()


if __name__ == "__main__":
    main()
