"""orbm_area_candidates_static (include/orbm.h): the window search whose scratch is a file-scope
`static __device__` array -- no cudaMalloc / cudaFree inside the call.  Checked through the C ABI against
the existing entry point orbm_area_candidates, bit for bit (row_ptr, candidate order, distances), on two
calls in a row with different inputs (the array keeps the first call's contents), plus the size limits.
Collected last on purpose (the file name): a failure here must not hide the other tests under `pytest -x`.
torch is used by the TEST only, to own the device buffers the entry point is given."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SF = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def _case(seed, n, nq, w=752, h=480):
    from orb_slam3_mac_b200._lib import KP_DTYPE
    from orb_slam3_mac_b200._lib_match import AREA_QUERY_DTYPE
    rng = np.random.default_rng(seed)
    keys = np.zeros(n, KP_DTYPE)
    keys["x"] = rng.uniform(-5, w + 5, n).astype(np.float32)     # a few keypoints fall outside the grid
    keys["y"] = rng.uniform(-5, h + 5, n).astype(np.float32)
    keys["x"][: n // 4] = np.round(keys["x"][: n // 4])
    keys["octave"] = rng.integers(0, 8, n)
    desc = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    q = np.zeros(nq, AREA_QUERY_DTYPE)
    q["x"] = rng.uniform(-40, w + 40, nq).astype(np.float32)
    q["y"] = rng.uniform(-40, h + 40, nq).astype(np.float32)
    q["r"] = rng.choice(np.array([0.0, 2.5, 7.0, 24.0, 100.0], np.float32), nq)
    gates = [(-1, -1), (0, 0), (2, 3), (1, -1), (0, 7), (5, 2)]
    for i in range(nq):
        q["min_level"][i], q["max_level"][i] = gates[i % len(gates)]
    q["valid"] = (rng.uniform(size=nq) > 0.1).astype(np.int32)
    qdesc = rng.integers(0, 256, (257, 32), dtype=np.uint8)
    q["desc_row"] = rng.integers(0, 257, nq)
    return keys, desc, q, qdesc, w, h


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()


def _static_call(lib, F, keys, desc, q, qdesc, cap):
    import torch
    d_keys, d_desc, d_q, d_qd = _dev(keys), _dev(desc), _dev(q), _dev(qdesc)
    d_rp = torch.full((len(q) + 1,), -7, dtype=torch.int32, device="cuda")
    d_cand = torch.full((max(cap, 1),), -7, dtype=torch.int32, device="cuda")
    d_dist = torch.zeros((max(cap, 1),), dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.orbm_area_candidates_static(
        torch.cuda.current_device(), C.c_void_p(stream), C.c_void_p(d_keys.data_ptr()), C.c_void_p(d_desc.data_ptr()),
        len(keys), F.bounds[0], F.bounds[1], F.grid_w_inv, F.grid_h_inv, C.c_void_p(d_q.data_ptr()), len(q),
        C.c_void_p(d_qd.data_ptr()), C.c_void_p(d_rp.data_ptr()), C.c_void_p(d_cand.data_ptr()),
        C.c_void_p(d_dist.data_ptr()), cap)
    assert rc == 0, rc
    torch.cuda.synchronize()
    rp = d_rp.cpu().numpy()
    nnz = int(rp[-1])
    return rp, d_cand.cpu().numpy()[:min(nnz, cap)], d_dist.cpu().numpy().view(np.uint16)[:min(nnz, cap)]


def test_static_scratch_window_search_matches_the_allocating_entry_point():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from orb_slam3_mac_b200 import _lib
    from orb_slam3_mac_b200.matcher import Frame, ORBmatcher
    lib = _lib.load()
    m = ORBmatcher(0.8, True)
    try:
        # two calls in a row: a large frame (> 4096 keypoints: the kernel's cell-of-keypoint scratch path) first,
        # then a smaller one with fewer queries, which must not see anything of the first call's grid or counts
        for seed, n, nq in ((1, 9000, 5000), (2, 1800, 1200)):
            keys, desc, q, qdesc, w, h = _case(seed, n, nq)
            F = Frame(keys, desc, w, h, SF)
            rp_ref, cand_ref, dist_ref = m.area_candidates(F, q, qdesc, cap=4_000_000)
            assert len(cand_ref) > 10000
            rp, cand, dist = _static_call(lib, F, keys, desc, q, qdesc, cap=len(cand_ref) + 64)
            assert np.array_equal(rp, rp_ref), (seed, "row_ptr")
            assert np.array_equal(cand, cand_ref), (seed, "candidates")       # bit-exact: integer work
            assert np.array_equal(dist, dist_ref), (seed, "distances")
        # a capacity below nnz: the count is still exact, the pairs that fit are the prefix
        rp, cand, dist = _static_call(lib, F, keys, desc, q, qdesc, cap=1000)
        assert np.array_equal(rp, rp_ref) and np.array_equal(cand, cand_ref[:1000]) and np.array_equal(dist, dist_ref[:1000])
        # beyond the stated maxima: a clear error, nothing launched
        from orb_slam3_mac_b200._lib_match import ORBM_STATIC_MAX_KEYS, ORBM_STATIC_MAX_QUERIES
        one = torch.zeros(64, dtype=torch.uint8, device="cuda")
        p = C.c_void_p(one.data_ptr())
        rc = lib.orbm_area_candidates_static(0, None, p, p, ORBM_STATIC_MAX_KEYS + 1, 0.0, 0.0, 1.0, 1.0, p, 1, p, p, p, p, 1)
        assert rc == _lib.ORBX_E_TOO_LARGE and b"ORBM_STATIC_MAX_KEYS" in lib.orbx_last_error_string()
        rc = lib.orbm_area_candidates_static(0, None, p, p, 1, 0.0, 0.0, 1.0, 1.0, p, ORBM_STATIC_MAX_QUERIES + 1, p, p, p, p, 1)
        assert rc == _lib.ORBX_E_TOO_LARGE and b"ORBM_STATIC_MAX_QUERIES" in lib.orbx_last_error_string()
    finally:
        m.close()
