"""Seeded synthetic inputs: no datasets are available offline.

Everything is numpy on the host; identical bytes feed the CPU oracle and the CUDA path.
"""
from __future__ import annotations

import numpy as np

SHAPES = {
    "C1": (480, 752),    # EuRoC mono, 1000 features
    "C2": (480, 752),    # EuRoC stereo, 1200 features per image
    "C3": (512, 512),    # TUM-VI fisheye stereo, 1500 features
    "C4": (720, 1280),   # phone mono, 2000 features
    "C5": (1080, 1920),  # batched stream, 4000 features
}
NFEATURES = {"C1": 1000, "C2": 1200, "C3": 1500, "C4": 2000, "C5": 4000}


def _value_noise(rng: np.random.Generator, h: int, w: int, cell: int) -> np.ndarray:
    gh, gw = h // cell + 2, w // cell + 2
    g = rng.uniform(-1.0, 1.0, size=(gh, gw)).astype(np.float32)
    ys = np.arange(h, dtype=np.float32) / cell
    xs = np.arange(w, dtype=np.float32) / cell
    y0 = np.floor(ys).astype(np.int64)
    x0 = np.floor(xs).astype(np.int64)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def textured_noise(h: int, w: int, seed: int) -> np.ndarray:
    """Three octaves of value noise (cells 64/16/4 px, amplitudes 60/40/25) + N(0,4) + 128."""
    rng = np.random.default_rng(seed)
    img = np.full((h, w), 128.0, dtype=np.float32)
    for cell, amp in ((64, 60.0), (16, 40.0), (4, 25.0)):
        img += amp * _value_noise(rng, h, w, cell)
    img += rng.normal(0.0, 4.0, size=(h, w)).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def checkerboard(h: int, w: int, seed: int) -> np.ndarray:
    """Homography-warped checkerboard rendered at 2x and box-downsampled, + N(0,2)."""
    rng = np.random.default_rng(seed + 7919)
    nx = int(rng.integers(8, 25))
    ny = max(6, int(round(nx * 0.75)))
    contrast = float(rng.uniform(40, 200))
    base = float(rng.uniform(contrast / 2 + 5, 250 - contrast / 2))
    # image corners -> board coordinates with jitter (<= 15 %)
    src = np.array([[0, 0], [2 * w, 0], [2 * w, 2 * h], [0, 2 * h]], dtype=np.float64)
    dst = np.array([[0, 0], [nx, 0], [nx, ny], [0, ny]], dtype=np.float64)
    dst += rng.uniform(-0.15, 0.15, size=(4, 2)) * np.array([nx, ny])
    A = []
    for (x, y), (u, v) in zip(src, dst):
        A.append([x, y, 1, 0, 0, 0, -u * x, -u * y, -u])
        A.append([0, 0, 0, x, y, 1, -v * x, -v * y, -v])
    _, _, vt = np.linalg.svd(np.asarray(A))
    H = vt[-1].reshape(3, 3)
    yy, xx = np.mgrid[0:2 * h, 0:2 * w].astype(np.float64)
    den = H[2, 0] * xx + H[2, 1] * yy + H[2, 2]
    u = (H[0, 0] * xx + H[0, 1] * yy + H[0, 2]) / den
    v = (H[1, 0] * xx + H[1, 1] * yy + H[1, 2]) / den
    chk = ((np.floor(u).astype(np.int64) + np.floor(v).astype(np.int64)) & 1).astype(np.float32)
    hi = (chk * contrast + (base - contrast / 2)).astype(np.float32)
    lo = hi.reshape(h, 2, w, 2).mean(axis=(1, 3))
    lo += rng.normal(0.0, 2.0, size=(h, w)).astype(np.float32)
    return np.clip(np.rint(lo), 0, 255).astype(np.uint8)


def low_texture(h: int, w: int, seed: int) -> np.ndarray:
    """Constant + N(0,3): forces the minThFAST fallback and near-empty levels."""
    rng = np.random.default_rng(seed + 104729)
    img = 128.0 + rng.normal(0.0, 3.0, size=(h, w))
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def black(h: int, w: int) -> np.ndarray:
    return np.zeros((h, w), dtype=np.uint8)


def stereo_pair(h: int, w: int, seed: int):
    """Rectified pair: right = left resampled with a planar disparity of 5..60 px + own noise."""
    rng = np.random.default_rng(seed + 15485863)
    wide = textured_noise(h, w + 64, seed).astype(np.float32)
    left = wide[:, :w]
    d0 = float(rng.uniform(5, 30))
    d1 = float(rng.uniform(0, 30))
    ys = np.arange(h, dtype=np.float32)[:, None]
    disp = d0 + d1 * ys / max(h - 1, 1)  # per-row disparity keeps rows aligned
    xs = np.arange(w, dtype=np.float32)[None, :] + disp
    x0 = np.floor(xs).astype(np.int64)
    f = xs - x0
    rows = np.arange(h)[:, None]
    right = wide[rows, x0] * (1 - f) + wide[rows, np.minimum(x0 + 1, w + 63)] * f
    right = right + rng.normal(0.0, 2.0, size=(h, w)).astype(np.float32)
    to8 = lambda a: np.clip(np.rint(a), 0, 255).astype(np.uint8)
    return to8(left), to8(right)


def fisheye_pair(n: int, seed: int):
    """Two n x n frames of the same scene with a 20 px shift (config C3)."""
    wide = textured_noise(n, n + 20, seed)
    rng = np.random.default_rng(seed + 32452843)
    right = wide[:, 20:].astype(np.float32) + rng.normal(0.0, 2.0, size=(n, n))
    return np.ascontiguousarray(wide[:, :n]), np.clip(np.rint(right), 0, 255).astype(np.uint8)


def frame(kind: str, h: int, w: int, seed: int) -> np.ndarray:
    if kind == "noise":
        return textured_noise(h, w, seed)
    if kind == "checker":
        return checkerboard(h, w, seed)
    if kind == "lowtex":
        return low_texture(h, w, seed)
    if kind == "black":
        return black(h, w)
    raise ValueError(kind)


# ---- matcher inputs ------------------------------------------------------
MAP_POINT_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_xr", "<f4"), ("view_cos", "<f4"),
                            ("depth", "<f4"), ("level", "<i4"), ("observations", "<i4"),
                            ("in_view", "u1"), ("bad", "u1"), ("_pad", "u1", (2,))])
LAST_OBS_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("inv_z", "<f4"), ("angle", "<f4"),
                           ("octave", "<i4"), ("observations", "<i4"), ("valid", "u1"), ("_pad", "u1", (3,))])


def flip_bits(desc: np.ndarray, p: float, rng) -> np.ndarray:
    bits = np.unpackbits(desc, axis=1)
    bits ^= (rng.random(bits.shape) < p).astype(np.uint8)
    return np.packbits(bits, axis=1)


def map_points_from_frame(keys, desc, width, height, total: int, seed: int, nlevels: int = 8):
    """Map points that re-observe a frame's keypoints (descriptor bits flipped w.p. 0.04, projection
    = keypoint + N(0, 1.5 px), predicted level = octave, viewCos ~ U(0.9, 1)), padded to `total` with
    uniformly random descriptors at uniformly random in-image positions."""
    rng = np.random.default_rng(seed + 49979687)
    n = min(len(keys), total)
    mps = np.zeros(total, MAP_POINT_DTYPE)
    md = np.zeros((total, 32), np.uint8)
    md[:n] = flip_bits(desc[:n], 0.04, rng)
    mps["proj_x"][:n] = keys["x"][:n] + rng.normal(0, 1.5, n)
    mps["proj_y"][:n] = keys["y"][:n] + rng.normal(0, 1.5, n)
    mps["level"][:n] = keys["octave"][:n]
    md[n:] = rng.integers(0, 256, (total - n, 32), dtype=np.uint8)
    mps["proj_x"][n:] = rng.uniform(0, width, total - n)
    mps["proj_y"][n:] = rng.uniform(0, height, total - n)
    mps["level"][n:] = rng.integers(0, nlevels, total - n)
    mps["view_cos"] = rng.uniform(0.9, 1.0, total)
    mps["depth"] = rng.uniform(0.5, 60.0, total)
    mps["proj_xr"] = mps["proj_x"] - rng.uniform(1.0, 40.0, total)
    mps["observations"] = rng.integers(0, 6, total)
    mps["in_view"] = (rng.random(total) < 0.9).astype(np.uint8)
    mps["bad"] = (rng.random(total) < 0.02).astype(np.uint8)
    perm = rng.permutation(total)  # acceptance is order dependent: interleave real and padding points
    return mps[perm].copy(), md[perm].copy()


def last_frame_obs(keys, desc, seed: int):
    """Observations of a last frame re-projected into the current one (3 px shift + noise)."""
    rng = np.random.default_rng(seed + 67867967)
    n = len(keys)
    lo = np.zeros(n, LAST_OBS_DTYPE)
    lo["u"] = keys["x"] + 3.0 + rng.normal(0, 1.0, n)
    lo["v"] = keys["y"] + rng.normal(0, 1.0, n)
    inv_z = rng.uniform(-0.02, 1.0, n).astype(np.float32)  # a few points behind the camera
    for _ in range(4):   # keep values that ARE float(1 / z) of a float depth z, as src/ORBmatcher.cc:2001 computes invzc
        inv_z = (1.0 / (1.0 / inv_z.astype(np.float64)).astype(np.float32).astype(np.float64)).astype(np.float32)
    lo["inv_z"] = inv_z
    lo["angle"] = keys["angle"]
    lo["octave"] = keys["octave"]
    lo["observations"] = rng.integers(0, 4, n)
    lo["valid"] = (rng.random(n) < 0.8).astype(np.uint8)
    return lo, flip_bits(desc, 0.03, rng)
