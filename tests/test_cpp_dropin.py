"""The C++ drop-in class (orb-slam3-mac_b200/host/ORBextractor.h, same shape as the reference's
include/ORBextractor.h:43-109) driven exactly like Frame::ExtractORB, compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp):
    exe = os.path.join(tmp, "dropin")
    pkg = os.path.join(ROOT, "orb-slam3-mac_b200")
    cmd = ["g++", "-std=c++11", "-O2", "-I" + os.path.join(pkg, "host"), "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "oracle", "cvshim"), os.path.join(ROOT, "tests", "cpp", "dropin_main.cpp"),
           "-L" + pkg, "-lorbx_b200", "-Wl,-rpath," + pkg, "-o", exe]
    subprocess.run(cmd, check=True)
    return exe


def test_dropin_compiles_against_the_cv_facade(tmp_path):
    """CPU check: the header has the reference's class shape and links against the C ABI."""
    _build(str(tmp_path))


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,nf,lap", [(480, 752, 1000, (0, 1000)), (512, 512, 1500, (0, 511)), (720, 1280, 2000, (0, 1000))])
def test_dropin_matches_oracle(tmp_path, oracle_libs, synth, h, w, nf, lap):
    from oracle.pyoracle import KP_DTYPE
    exe = _build(str(tmp_path))
    img = synth.textured_noise(h, w, 17)
    raw, out = str(tmp_path / "in.raw"), str(tmp_path / "out.bin")
    img.tofile(raw)
    level = 3
    subprocess.run([exe, str(w), str(h), str(nf), str(lap[0]), str(lap[1]), str(level), raw, out], check=True)
    b = open(out, "rb").read()
    hdr = np.frombuffer(b, np.int32, 6)
    mono, n, empty, nlev, lw, lh = [int(x) for x in hdr]
    off = 24
    sf = np.frombuffer(b, np.float32, 8, off); off += 32
    is2 = np.frombuffer(b, np.float32, 8, off); off += 32
    k = np.frombuffer(b, KP_DTYPE, n, off); off += 28 * n
    d = np.frombuffer(b, np.uint8, 32 * n, off).reshape(n, 32); off += 32 * n
    pyr = np.frombuffer(b, np.uint8, (lw + 38) * (lh + 38), off).reshape(lh + 38, lw + 38)

    kind = "ref" if oracle_libs.have_ref() else "port"
    ora = oracle_libs.OracleExtractor(nf, 1.2, 8, 20, 7, kind=kind)
    mono_r, k_r, d_r = ora(img, lap)
    assert empty == -1 and nlev == 8
    assert (mono, n) == (mono_r, len(k_r))
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(k[f], k_r[f]), f
    assert np.abs(k["angle"] - k_r["angle"]).max() <= 1e-4 * 180 / np.pi
    assert np.unpackbits(d ^ d_r).sum() <= 1e-3 * d.size * 8
    t = ora.tables()
    assert np.array_equal(sf, t["scale"]) and np.array_equal(is2, t["inv_sigma2"])
    assert np.array_equal(pyr, ora.level(level, 19))  # mvImagePyramid incl. the REFLECT_101 border
