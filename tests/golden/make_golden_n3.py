"""Freezes what the REFERENCE'S OWN code (src/ORBmatcher.cc, src/MapPoint.cc, DBoW2, compiled verbatim:
MatchOracle("ref")) computes on the N3 cases (tests/golden/n3_cases.py) into n3_golden.json; our restatement
(MatchOracle("port")) must agree before anything is written.
usage: python -m tests.golden.make_golden_n3"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import pyoracle as po  # noqa: E402
from tests.golden import n3_cases as nc  # noqa: E402

if __name__ == "__main__":
    po.build("mref")
    res = nc.digest(nc.run_n3_cases(po.MatchOracle("ref")))
    assert res == nc.digest(nc.run_n3_cases(po.MatchOracle("port"))), "port != reference"
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "n3_golden.json")
    json.dump(res, open(path, "w"), indent=1, sort_keys=True)
    for k, v in res.items():
        print(k, {a: b for a, b in v.items() if isinstance(b, int)})
