"""Developer probe: where the end-to-end time of a small layer-2 search goes (fixed per-call cost of the
copy + launch + sync sequence against payload size).  Prints median microseconds."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from orb_slam3_mac_b200.matcher import ORBmatcher


def med(f, n=200):
    for _ in range(20):
        f()
    t = []
    for _ in range(n):
        a = time.perf_counter()
        f()
        t.append(time.perf_counter() - a)
    return round(1e6 * float(np.median(t)), 1)


m = ORBmatcher(0.8, True)
rng = np.random.default_rng(0)
for nq, per in ((1, 1), (2004, 1), (2004, 22), (2004, 100)):
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (2005, 32), dtype=np.uint8)
    rp = (np.arange(nq + 1) * per).astype(np.int32)
    cand = rng.integers(0, 2005, nq * per).astype(np.int32)
    print("hamming_csr", nq, "rows x", per, "cands:", med(lambda: m.hamming_csr(q, t, rp, cand)), "us")
for nq in (1, 2004):
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (2005, 32), dtype=np.uint8)
    print("knn2", nq, "x 2005:", med(lambda: m.knn2(q, t)), "us")
