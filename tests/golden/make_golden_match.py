#!/usr/bin/env python
"""Generates tests/golden/match_golden.json (run in the build container):
  * matcher cases (match_cases.run_matcher_cases) from the REFERENCE'S OWN src/ORBmatcher.cc / Frame.cc,
    compiled verbatim (oracle/_ref/liborbmref.so, MatchOracle("ref")); our restatement (MatchOracle("port"))
    must agree before anything is written;
  * the input-stage case from the REAL OpenCV of this image (cv2.remap + cv2.cvtColor)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

if __name__ == "__main__":
    import cv2
    import orb_slam3_mac_b200  # noqa: F401
    from oracle import pyoracle as po
    from tests.golden import match_cases as mc
    po.build("port")
    po.build("mref")
    out = {"matcher": mc.run_matcher_cases(po.MatchOracle("ref")), "source": "reference src/ORBmatcher.cc, verbatim"}
    assert out["matcher"] == mc.run_matcher_cases(po.MatchOracle("port")), "port != reference"
    cv2.setNumThreads(1)
    bgr, mx, my = mc.input_stage_case()
    rect = cv2.remap(bgr, mx, my, cv2.INTER_LINEAR)
    out["input_stage"] = {"cv2": cv2.__version__, "bgr_rect_gray": mc.sha(cv2.cvtColor(rect, cv2.COLOR_BGR2GRAY)),
                          "rect_gray_only": mc.sha(cv2.remap(bgr[:, :, 1].copy(), mx, my, cv2.INTER_LINEAR)),
                          "rgb_gray": mc.sha(cv2.cvtColor(bgr, cv2.COLOR_RGB2GRAY))}
    json.dump(out, open(os.path.join(os.path.dirname(__file__), "match_golden.json"), "w"), indent=1)
    print(json.dumps(out, indent=1)[:1500])
