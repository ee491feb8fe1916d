"""Pins oracle/cvprims.h (the restated OpenCV primitives) bit-for-bit against the real OpenCV
in this image (cv2 4.13).  SURVEY.md 8(c): the reference holds no golden vectors for this
path; the arithmetic lives in un-vendored OpenCV, so cv2 is the witness."""
import ctypes as C

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


@pytest.fixture(scope="module")
def lib(oracle_libs):
    cv2.setNumThreads(1)
    l = C.CDLL(oracle_libs.PORT_SO)
    l.orbo_prim_fast.restype = C.c_int
    l.orbo_prim_round.restype = C.c_int
    l.orbo_prim_round.argtypes = [C.c_float]
    return l


def _p(a):
    return C.c_void_p(a.ctypes.data)


@pytest.mark.parametrize("shape,dshape", [
    ((480, 752), (400, 627)), ((400, 627), (333, 522)), ((512, 512), (427, 427)),
    ((720, 1280), (600, 1067)), ((1080, 1920), (900, 1600)), ((134, 210), (112, 175)),
    ((97, 131), (80, 109)), ((60, 75), (61, 90)), ((40, 40), (100, 93))])
def test_resize_matches_cv2(lib, shape, dshape):
    rng = np.random.default_rng(shape[0] * 7 + dshape[1])
    src = rng.integers(0, 256, size=shape, dtype=np.uint8)
    dst = np.zeros(dshape, np.uint8)
    lib.orbo_prim_resize(_p(src), shape[0], shape[1], C.c_size_t(src.strides[0]), _p(dst),
                         dshape[0], dshape[1], C.c_size_t(dst.strides[0]))
    ref = cv2.resize(src, (dshape[1], dshape[0]), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(dst, ref)


def test_resize_chain_matches_cv2(lib, synth):
    """The reference's pyramid: every level is resized from the previous ROUNDED level."""
    img = synth.textured_noise(480, 752, 3)
    cur = img
    for l in range(1, 8):
        s = np.float32(1.0) / np.float32(1.2) ** l
        w = int(np.rint(np.float32(752) * np.float32(1.0 / (1.2 ** l))))
        h = int(np.rint(np.float32(480) * np.float32(1.0 / (1.2 ** l))))
        dst = np.zeros((h, w), np.uint8)
        lib.orbo_prim_resize(_p(cur), cur.shape[0], cur.shape[1], C.c_size_t(cur.strides[0]),
                             _p(dst), h, w, C.c_size_t(dst.strides[0]))
        ref = cv2.resize(cur, (w, h), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(dst, ref), l
        cur = ref


@pytest.mark.parametrize("shape", [(480, 752), (512, 512), (134, 210), (33, 47), (9, 200), (7, 7)])
def test_blur_matches_cv2(lib, shape, synth):
    for seed, gen in enumerate((lambda: np.random.default_rng(1).integers(0, 256, shape, dtype=np.uint8),
                                lambda: synth.textured_noise(shape[0], shape[1], 5))):
        src = gen()
        dst = np.zeros_like(src)
        lib.orbo_prim_blur(_p(src), shape[0], shape[1], C.c_size_t(src.strides[0]), _p(dst),
                           C.c_size_t(dst.strides[0]))
        ref = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(dst, ref)


def test_border_matches_cv2(lib):
    src = np.random.default_rng(2).integers(0, 256, (40, 61), dtype=np.uint8)
    dst = np.zeros((40 + 38, 61 + 38), np.uint8)
    lib.orbo_prim_border(_p(src), 40, 61, C.c_size_t(src.strides[0]), _p(dst),
                         C.c_size_t(dst.strides[0]), 19)
    ref = cv2.copyMakeBorder(src, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    assert np.array_equal(dst, ref)


def _cv_fast(img, th):
    det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    kps = det.detect(np.ascontiguousarray(img))
    for k in kps:
        assert k.size == 7 and k.angle == -1 and k.octave == 0 and k.class_id == -1
    return np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)


def test_fast_matches_cv2(lib, synth):
    rng = np.random.default_rng(11)
    total = 0
    big = synth.textured_noise(480, 752, 9)
    chk = synth.checkerboard(480, 752, 4)
    for it in range(120):
        h, w = int(rng.integers(7, 46)), int(rng.integers(7, 46))
        if it % 3 == 0:
            img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        else:
            srcimg = big if it % 3 == 1 else chk
            y0, x0 = int(rng.integers(0, 480 - h)), int(rng.integers(0, 752 - w))
            img = srcimg[y0:y0 + h, x0:x0 + w]  # strided view, like the reference's cell views
        for th in (20, 7):
            out = np.zeros((h * w, 3), np.int32)
            n = lib.orbo_prim_fast(_p(img), h, w, C.c_size_t(img.strides[0]), th, _p(out), h * w)
            ref = _cv_fast(img, th)
            assert n == len(ref)
            assert np.array_equal(out[:n], ref)
            total += n
    assert total > 1000


def test_fast_whole_image_matches_cv2(lib, synth):
    img = synth.textured_noise(240, 376, 13)
    out = np.zeros((img.size, 3), np.int32)
    n = lib.orbo_prim_fast(_p(img), 240, 376, C.c_size_t(img.strides[0]), 20, _p(out), img.size)
    ref = _cv_fast(img, 20)
    assert n == len(ref) and np.array_equal(out[:n], ref)


def test_fast_atan2_matches_cv2(lib):
    rng = np.random.default_rng(5)
    y = rng.integers(-2_700_000, 2_700_000, 200_000).astype(np.float32)
    x = rng.integers(-2_700_000, 2_700_000, 200_000).astype(np.float32)
    y[:1000] = 0
    x[500:1500] = 0
    out = np.zeros_like(x)
    lib.orbo_prim_atan2(_p(y), _p(x), _p(out), len(x))
    # cv2.fastAtan2 is the scalar function the reference calls (src/ORBextractor.cc:101);
    # cv2.phase() runs a differently-contracted SIMD variant and is deliberately not used.
    ref = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert np.array_equal(out, ref)


def test_cv_round_is_half_even(lib):
    for v, r in ((0.5, 0), (1.5, 2), (2.5, 2), (-0.5, 0), (-1.5, -2), (2.4999, 2), (3.5, 4)):
        assert lib.orbo_prim_round(v) == r


@pytest.mark.parametrize("code,channels,rgb", [("COLOR_BGR2GRAY", 3, 0), ("COLOR_RGB2GRAY", 3, 1),
                                               ("COLOR_BGRA2GRAY", 4, 0), ("COLOR_RGBA2GRAY", 4, 1)])
def test_cvt_gray_matches_cv2(lib, code, channels, rgb):
    """N4: cvtColor to gray (src/Tracking.cc:354-379).  cv2 4.13 uses the 15-bit coefficients."""
    rng = np.random.default_rng(channels * 2 + rgb)
    for shape in [(480, 752), (37, 53), (1, 1)]:
        src = rng.integers(0, 256, size=shape + (channels,), dtype=np.uint8)
        dst = np.zeros(shape, np.uint8)
        lib.orbo_prim_cvt_gray(_p(src), shape[0], shape[1], C.c_size_t(src.strides[0]), channels, rgb, 15,
                               _p(dst), C.c_size_t(dst.strides[0]))
        assert np.array_equal(dst, cv2.cvtColor(src, getattr(cv2, code)))
    # all 2^24 colours once
    a = np.arange(1 << 24, dtype=np.uint32)
    src = np.stack([a & 255, (a >> 8) & 255, a >> 16], axis=1).astype(np.uint8).reshape(4096, 4096, 3)
    dst = np.zeros((4096, 4096), np.uint8)
    lib.orbo_prim_cvt_gray(_p(src), 4096, 4096, C.c_size_t(src.strides[0]), 3, 0, 15, _p(dst),
                           C.c_size_t(dst.strides[0]))
    assert np.array_equal(dst, cv2.cvtColor(src, cv2.COLOR_BGR2GRAY))


def _euroc_maps():
    K = np.array([[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]])
    D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0])
    R = np.array([[0.999966347530033, -0.001422739138722922, 0.008079580483432283],
                  [0.001365741834644127, 0.9999741760894847, 0.007055629199258132],
                  [-0.008089410156878961, -0.007044357138835809, 0.9999424675829176]])
    P = np.array([[435.2046959714599, 0, 367.4517211914062], [0, 435.2046959714599, 252.2008514404297],
                  [0, 0, 1]])
    return cv2.initUndistortRectifyMap(K, D, R, P, (752, 480), cv2.CV_32F)


@pytest.mark.parametrize("channels", [1, 3, 4])
def test_remap_linear_matches_cv2(lib, channels):
    """N4: cv::remap(INTER_LINEAR, BORDER_CONSTANT 0) with CV_32FC1 maps
    (Examples/Stereo/stereo_euroc.cc:170-171): EuRoC rectification maps, noisy maps that leave the
    image on every side, and exact half-way fractions (cvRound ties)."""
    rng = np.random.default_rng(40 + channels)
    cases = []
    m1, m2 = _euroc_maps()
    cases.append(((480, 752), m1, m2))
    for (h, w), amp in [((120, 160), 0.5), ((120, 160), 60.0), ((33, 47), 8.0)]:
        yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
        mx = (xx + 3.3 * np.sin(yy / 17) + rng.uniform(-amp, amp, (h, w))).astype(np.float32)
        my = (yy + 2.7 * np.cos(xx / 23) + rng.uniform(-amp, amp, (h, w))).astype(np.float32)
        cases.append(((h, w), mx, my))
    yy, xx = np.mgrid[0:64, 0:64].astype(np.float32)
    cases.append(((64, 64), xx + np.float32(1 / 64), yy + np.float32(3 / 64)))
    cases.append(((64, 64), xx - np.float32(1 / 64) - 5, yy + np.float32(63 / 64)))
    for (h, w), mx, my in cases:
        shape = (h, w) if channels == 1 else (h, w, channels)
        src = rng.integers(0, 256, size=shape, dtype=np.uint8)
        mx, my = np.ascontiguousarray(mx), np.ascontiguousarray(my)
        dst = np.zeros(shape, np.uint8)
        lib.orbo_prim_remap(_p(src), h, w, C.c_size_t(src.strides[0]), channels, _p(mx), _p(my), _p(dst), h, w,
                            C.c_size_t(dst.strides[0]))
        assert np.array_equal(dst, cv2.remap(src, mx, my, cv2.INTER_LINEAR)), (h, w)
