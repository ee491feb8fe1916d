#!/usr/bin/env python
"""Generates tests/golden/extractor_golden.json from the reference's own ORBextractor.cc
(oracle/_ref, built from /root/reference -- only possible in the build container).

For each fixed-seed synthetic frame: sha1 of every pyramid level, of the per-level FAST candidate
lists, of the per-level quadtree survivors, and of the final keypoint / descriptor arrays.
The port (oracle/orb_oracle.cpp) and the CUDA path are tested against these hashes where
/root/reference (hence oracle/_ref) is not available.
"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

CASES = [
    ("noise", 480, 752, 1000, (0, 1000), 0), ("noise", 480, 752, 1200, (0, 0), 1),
    ("noise", 512, 512, 1500, (0, 511), 2), ("noise", 720, 1280, 2000, (0, 1000), 3),
    ("noise", 1080, 1920, 4000, (0, 0), 4), ("noise", 480, 752, 5000, (0, 1000), 5),
    ("checker", 480, 752, 1000, (0, 1000), 6), ("lowtex", 480, 752, 1000, (0, 0), 7),
    ("black", 480, 752, 1000, (0, 0), 8), ("noise", 333, 517, 700, (100, 300), 9),
    ("checker", 240, 1000, 500, (0, 0), 10), ("noise", 600, 400, 800, (0, 0), 11),
]


def sha(a):
    return hashlib.sha1(a.tobytes()).hexdigest()


def digest(ex, img, lap):
    mono, k, d = ex(img, lap)
    cands, survs = ex.stages(img)
    return {
        "mono": int(mono), "n": int(len(k)), "kps": sha(k), "desc": sha(d),
        "levels": [sha(ex.level(l)) for l in range(8)],
        "cands": [sha(c) for c in cands], "n_cands": [int(len(c)) for c in cands],
        "survs": [sha(s) for s in survs], "n_survs": [int(len(s)) for s in survs],
    }


if __name__ == "__main__":
    import orb_slam3_mac_b200 as pkg
    from oracle import pyoracle as po
    po.build("ref")
    assert po.have_ref(), "needs /root/reference"
    out = {}
    for kind, h, w, nf, lap, seed in CASES:
        ex = po.OracleExtractor(nf, 1.2, 8, 20, 7, kind="ref")
        out[f"{kind}-{h}x{w}-{nf}-{lap[0]}_{lap[1]}-s{seed}"] = digest(ex, pkg.synth.frame(kind, h, w, seed), lap)
    json.dump(out, open(os.path.join(os.path.dirname(__file__), "extractor_golden.json"), "w"), indent=1)
    print("wrote", len(out), "cases")
