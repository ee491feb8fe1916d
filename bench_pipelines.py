#!/usr/bin/env python
"""The composite BASELINE.json configs as PIPELINES (extraction + the matcher step the config names), each beside
the reference's CPU code for the same pipeline on this box's host cores.  bench.py puts the result under
`configs` in its JSON line (outside the timed C1 region); run alone it prints one JSON line per config.

  C2  EuRoC stereo pair 2 x 752x480, 1200 features: left and right extraction on two handles / two host threads
      (src/Frame.cc:109-112) + Frame::ComputeStereoMatches on the two DEVICE-RESIDENT pyramids (:802-980)
  C3  fisheye stereo 2 x 512x512, 1500 features, lapping {0, 511}: two extractions + BFMatcher.knnMatch k=2 and
      the 0.7 ratio test (:1056-1059, :1128-1153)
  C4  1280x720, 2000 features + SearchByProjection against 10,000 map points (src/ORBmatcher.cc:48-147), the
      frame kept on the device between extraction and search (orbm_frame_from_device)
Every unit goes host image(s) in -> host result out, synchronously; p50 over `reps` units, value = 1 / p50.
CPU arm: the reference's own ORBextractor.cc (verbatim, oracle/_ref) with the reference's threading for the
pair, and the matcher loops of oracle/match_oracle.cpp (our port, pinned bit-for-bit to the verbatim build of
src/ORBmatcher.cc / Frame.cc by tests/) on one thread.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def _p50(f, reps, warm=3):
    for _ in range(warm):
        f(0)
    ts = []
    for i in range(reps):
        t0 = time.perf_counter()
        f(i)
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e6


def run_pipelines(reps: int = 40, cpu: bool = True, device: int = 0) -> dict:
    import orb_slam3_mac_b200 as pkg  # noqa: F401
    from orb_slam3_mac_b200 import _lib, synth as S
    from orb_slam3_mac_b200.extractor import ORBextractor
    from orb_slam3_mac_b200.matcher import Frame, ORBmatcher, ResidentFrame
    po = mo = None
    if cpu:
        from oracle import pyoracle as po
        # timed: our port of the reference loops (the verbatim build's wrapper constructs Frame / MapPoint objects per
        # call, which the reference does not pay); the verbatim build is the parity checker of tests/
        mo = po.MatchOracle("port")
    kind = None if po is None else ("ref" if po.have_ref() else "port")
    out = {}
    pool = ThreadPoolExecutor(2)
    m = ORBmatcher(0.8, True, device)
    lib = m.lib

    # ---- C2 ---------------------------------------------------------------------------------------------
    pairs = [S.stereo_pair(480, 752, s) for s in range(4)]
    exL, exR = ORBextractor(1200, 1.2, 8, 20, 7, max_size=(480, 752)), ORBextractor(1200, 1.2, 8, 20, 7, max_size=(480, 752))
    mb, mbf = 0.11, 47.9
    res = {}

    def c2(i):
        l, r = pairs[i % len(pairs)]
        fl, fr = pool.submit(exL, l, (0, 0)), pool.submit(exR, r, (0, 0))
        (_, kl, dl), (_, kr, dr) = fl.result(), fr.result()
        L, R = Frame(kl, dl, 752, 480, SF), Frame(kr, dr, 752, 480, SF)
        res["u"], res["d"] = m.ComputeStereoMatches(L, R, exL.device_pyramid(), exR.device_pyramid(), mb, mbf)
        res["n"] = len(kl)

    g = _p50(c2, reps)
    row = {"what": "2 x ORB extraction (two handles, two host threads) + ComputeStereoMatches on the device-resident pyramids",
           "unit": "stereo pairs/s", "p50_us": g, "value": 1e6 / g, "keypoints_left": res["n"],
           "stereo_matches": int((res["u"] >= 0).sum())}
    if cpu:
        ol, orr = po.OracleExtractor(1200, 1.2, 8, 20, 7, kind=kind), po.OracleExtractor(1200, 1.2, 8, 20, 7, kind=kind)

        def c2cpu(i):
            l, r = pairs[i % len(pairs)]
            fl, fr = pool.submit(ol, l, (0, 0)), pool.submit(orr, r, (0, 0))
            (_, kl, dl), (_, kr, dr) = fl.result(), fr.result()
            L, R = Frame(kl, dl, 752, 480, SF), Frame(kr, dr, 752, 480, SF)
            pl = [ol.level(k, 19) for k in range(8)]
            pr = [orr.level(k, 19) for k in range(8)]
            res["cu"] = mo.compute_stereo_matches(L, R, pl, pr, mb, mbf)[0]

        c = _p50(c2cpu, max(4, reps // 8), warm=1)
        c2(0)
        c2cpu(0)
        assert np.array_equal(res["cu"].view(np.uint32), res["u"].view(np.uint32)), "C2 pipeline: GPU != reference"
        row.update({"cpu_p50_us": c, "cpu_threads": 2, "cpu_kind": f"extractor {kind}, matcher {mo.kind}", "speedup": c / g,
                    "parity": "mvuRight bit-identical to the CPU pipeline on one pair"})
    out["C2"] = row
    exL.close()
    exR.close()

    # ---- C3 ---------------------------------------------------------------------------------------------
    fpairs = [S.fisheye_pair(512, s) for s in range(4)]
    exL, exR = ORBextractor(1500, 1.2, 8, 20, 7, max_size=(512, 512)), ORBextractor(1500, 1.2, 8, 20, 7, max_size=(512, 512))

    def c3(i):
        l, r = fpairs[i % len(fpairs)]
        fl, fr = pool.submit(exL, l, (0, 511)), pool.submit(exR, r, (0, 511))
        (ml, kl, dl), (mr, kr, dr) = fl.result(), fr.result()
        idx, dist = m.knn2(dl[ml:], dr[mr:])
        res["good"] = (dist[:, 0] < dist[:, 1] * np.float32(0.7)) & (idx[:, 1] >= 0)
        res["idx"] = idx

    g = _p50(c3, reps)
    row = {"what": "2 x ORB extraction (lapping {0,511}) + brute-force 2-NN Hamming + 0.7 ratio test",
           "unit": "stereo pairs/s", "p50_us": g, "value": 1e6 / g, "ratio_matches": int(res["good"].sum())}
    if cpu:
        ol, orr = po.OracleExtractor(1500, 1.2, 8, 20, 7, kind=kind), po.OracleExtractor(1500, 1.2, 8, 20, 7, kind=kind)

        def c3cpu(i):
            l, r = fpairs[i % len(fpairs)]
            fl, fr = pool.submit(ol, l, (0, 511)), pool.submit(orr, r, (0, 511))
            (ml, kl, dl), (mr, kr, dr) = fl.result(), fr.result()
            res["cidx"] = mo.knn2(dl[ml:], dr[mr:])[0]

        c = _p50(c3cpu, max(4, reps // 8), warm=1)
        c3(0)
        c3cpu(0)
        assert np.array_equal(res["cidx"], res["idx"]), "C3 pipeline: GPU != reference"
        row.update({"cpu_p50_us": c, "cpu_threads": 2, "cpu_kind": f"extractor {kind}, 2-NN: scalar popcount loop",
                    "speedup": c / g, "parity": "2-NN indices identical to the CPU pipeline on one pair"})
    out["C3"] = row
    exL.close()
    exR.close()

    # ---- C4 ---------------------------------------------------------------------------------------------
    img0 = S.textured_noise(720, 1280, 7)
    imgs = [np.roll(img0, 3 + s, axis=1) for s in range(4)]
    ex = ORBextractor(2000, 1.2, 8, 20, 7, max_size=(720, 1280))
    _, k0, d0 = ex(img0, (0, 1000))
    mps, md = S.map_points_from_frame(k0, d0, 1280, 720, 10000, 7)
    hdl = ex.h
    cap = lib.orbx_max_keypoints(hdl)
    import torch
    d_imgs = [torch.from_numpy(im).cuda(device) for im in imgs]
    torch.cuda.synchronize()
    lap = np.array([[0, 1000]], np.int32)
    d_hdr, d_kps, d_desc = C.c_void_p(), C.c_void_p(), C.c_void_p()
    hdr = np.zeros(4, np.int32)

    def c4(i):
        im = imgs[i % len(imgs)]
        # host image in -> keypoints / descriptors to the host -> the search uploads the frame again (the round-1 path)
        _, k, d = ex(im, (0, 1000))
        F = Frame(k, d, 1280, 720, SF)
        res["n"], res["match"], _ = m.SearchByProjection(F, mps, md, 1.0, False, 30.0)
        res["nk"] = len(k)

    g = _p50(c4, reps)
    row = {"what": "ORB extraction (2000 features) + SearchByProjection of 10,000 map points, host image in / matches out",
           "unit": "frames/s", "p50_us": g, "value": 1e6 / g, "keypoints": res["nk"], "matches": int(res["n"])}

    def c4dev(i):
        di = d_imgs[i % len(d_imgs)]
        _lib.check(lib.orbx_extract_device(hdl, 0, di.data_ptr(), 1, 1280, 720, 1280, 1280 * 720, lap.ctypes.data,
                                           C.byref(d_hdr), C.byref(d_kps), C.byref(d_desc)))
        _lib.check(lib.orbx_wait(hdl, 0))
        n = int(torch.as_tensor(_Dev(d_hdr.value, 16), device=f"cuda:{device}").cpu().numpy().view(np.int32)[0])
        RF = ResidentFrame(m, d_keys=d_kps.value, d_desc=d_desc.value, n=n, camera=None, width=1280, height=720, scale_factors=SF)
        res["n_dev"], res["match_dev"], _ = m.SearchByProjection(RF, mps, md, 1.0, False, 30.0)
        RF.close()

    gd = _p50(c4dev, reps)
    row["device_resident"] = {"what": "frame already in HBM (orbx_extract_device) -> orbm_frame_from_device -> search: keypoints and "
                                      "descriptors never visit the host on the way", "p50_us": gd, "value": 1e6 / gd}
    if cpu:
        o = po.OracleExtractor(2000, 1.2, 8, 20, 7, kind=kind)

        def c4cpu(i):
            im = imgs[i % len(imgs)]
            _, k, d = o(im, (0, 1000))
            F = Frame(k, d, 1280, 720, SF)
            res["cn"], res["cmatch"], _ = mo.search_by_projection_points(F, mps, md, 1.0, False, 30.0, 0.8)

        c = _p50(c4cpu, max(4, reps // 8), warm=1)
        c4(0)
        c4cpu(0)
        assert res["cn"] == res["n"] and np.array_equal(res["cmatch"], res["match"]), "C4 pipeline: GPU != reference"
        row.update({"cpu_p50_us": c, "cpu_threads": 1, "cpu_kind": f"extractor {kind}, matcher {mo.kind}", "speedup": c / g,
                    "parity": "match indices identical to the CPU pipeline on one frame"})
    out["C4"] = row
    ex.close()
    m.close()
    pool.shutdown()
    return out


class _Dev:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


if __name__ == "__main__":
    for k, v in run_pipelines().items():
        print(json.dumps({"config": k, **v}), flush=True)
