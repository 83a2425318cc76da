"""numpy restatements shared by the generator / dataset tests (test infrastructure, not product code):
the Q14 nearest-neighbour rotation and the max-overlap label search of magic_b200/csrc/kernels.cu in integer
arithmetic (bit-exact), a float32 restatement of tf.contrib.image.rotate's inverse map (8c/utils.py:329-340), the
first layer's operand-stage density of a batch, and batches sampled from the reference's real images
(tests/golden/overlap_micro_100.npz)."""
from __future__ import annotations

import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROT_CS = [16384, 11585, 0, -11585]      # Q14 cos/sin of r * 45 degrees, r = 0..3 (kernels.cu: rot_cs / rot_sn)
ROT_SN = [0, 11585, 16384, 11585]


def _grid(side):
    return np.meshgrid(np.arange(side, dtype=np.int64), np.arange(side, dtype=np.int64), indexing="ij")


def nn_rotate(img: np.ndarray, cs: int, sn: int) -> np.ndarray:
    """Nearest-neighbour inverse-map rotation about the canvas centre in Q14 fixed point on doubled coordinates
    (kernels.cu: rotated_index), zero fill."""
    side = img.shape[0]
    y, x = _grid(side)
    X2, Y2 = 2 * x - side + 1, 2 * y - side + 1
    sx = (cs * X2 - sn * Y2 + (side << 14)) >> 15
    sy = (sn * X2 + cs * Y2 + (side << 14)) >> 15
    ok = (sx >= 0) & (sx < side) & (sy >= 0) & (sy < side)
    out = np.zeros_like(img)
    out[ok] = img[sy[ok], sx[ok]]
    return out


def tf_rotate_float(img: np.ndarray, angle: float) -> np.ndarray:
    """tf.contrib.image.rotate(img, angle) with the default NEAREST interpolation, restated in float32 from the TF 1.x
    sources (third-party, absent from /root/reference; called at 8c/utils.py:337-338):
    angles_to_projective_transforms builds  [cos, -sin, x_off, sin, cos, y_off, 0, 0]  with
    x_off = ((W-1) - (cos (W-1) - sin (H-1))) / 2,  y_off = ((H-1) - (sin (W-1) + cos (H-1))) / 2  and the kernel reads
    input(round(a0 x + a1 y + a2), round(b0 x + b1 y + b2)) for output pixel (x, y), zero outside."""
    h, w = img.shape
    f = np.float32
    c, s = f(np.cos(f(angle))), f(np.sin(f(angle)))
    x_off = ((f(w) - f(1)) - (c * (f(w) - f(1)) - s * (f(h) - f(1)))) / f(2)
    y_off = ((f(h) - f(1)) - (s * (f(w) - f(1)) + c * (f(h) - f(1)))) / f(2)
    y, x = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    sx = np.rint(c * x - s * y + x_off).astype(np.int64)      # std::round differs from rint only at exact .5
    sy = np.rint(s * x + c * y + y_off).astype(np.int64)
    ok = (sx >= 0) & (sx < w) & (sy >= 0) & (sy < h)
    out = np.zeros_like(img)
    out[ok] = img[sy[ok], sx[ok]]
    return out


def centroid(img: np.ndarray):
    """Rounded pixel centroid (cx, cy) = (2 sum + area) // (2 area), area clamped to >= 1 (kernels.cu, phase 3a)."""
    ys, xs = np.nonzero(img)
    a = max(len(xs), 1)
    return int((2 * int(xs.sum()) + a) // (2 * a)), int((2 * int(ys.sum()) + a) // (2 * a))


def aligned_overlap(lock: np.ndarray, key: np.ndarray, r: int, tx: int = 0, ty: int = 0, stride: int = 1) -> int:
    """popcount(lock AND T(key)) for rotation r (0..7) and shift (tx, ty): lock pixel (x, y) reads key pixel
    cK +- ((cs X - sn Y + 8192) >> 14, (sn X + cs Y + 8192) >> 14), X = x - cLx - tx, Y = y - cLy - ty.
    stride > 1: only lock pixels with x % stride == 0 and y % stride == 0 are counted (the search grid)."""
    side = lock.shape[0]
    (clx, cly), (ckx, cky) = centroid(lock), centroid(key)
    cs, sn, sgn = ROT_CS[r & 3], ROT_SN[r & 3], (-1 if r >= 4 else 1)
    y, x = _grid(side)
    X, Y = x - clx - tx, y - cly - ty
    sx = ckx + sgn * ((cs * X - sn * Y + 8192) >> 14)
    sy = cky + sgn * ((sn * X + cs * Y + 8192) >> 14)
    ok = (sx >= 0) & (sx < side) & (sy >= 0) & (sy < side) & (lock > 0) & (x % stride == 0) & (y % stride == 0)
    return int((key[sy[ok], sx[ok]] > 0).sum())


def max_overlap_label(lock: np.ndarray, key: np.ndarray) -> int:
    """The generator's label (kernels.cu, phase 3; README.md:19-27 "maximum possible overlap area").  Search on every
    4th pixel: 8 rotations with the centroids aligned, then the first best rotation shifted by +-s along x and y,
    s = (3 side + 50) // 100; the label is the exact overlap of the first best of those five candidates."""
    side = lock.shape[0]
    stage1 = [aligned_overlap(lock, key, r, stride=4) for r in range(8)]
    rb = int(np.argmax(stage1))                              # first maximum, like the kernel's strict '>' scan
    s = (3 * side + 50) // 100
    shifts = [(0, 0), (s, 0), (-s, 0), (0, s), (0, -s)]
    stage2 = [stage1[rb]] + [aligned_overlap(lock, key, rb, tx, ty, stride=4) for tx, ty in shifts[1:]]
    tx, ty = shifts[int(np.argmax(stage2))]
    return aligned_overlap(lock, key, rb, tx, ty)


def stage_density(planes: np.ndarray, B: int):
    """Fraction of the first layer's 128 x 64 operand stages that hold a set pixel, for rows [rot; lock; key] of
    `planes` [3B, D] (bool): (forward GEMM, weight-gradient GEMM) -- the recount of magic_input_density."""
    D = planes.shape[1]
    kb = (D + 1 + 63) // 64
    rows = (3 * B + 127) // 128 * 128
    P = np.zeros((rows, kb * 64), dtype=bool)
    P[:3 * B, :D] = planes
    P[:3 * B, D] = True                                      # bias column
    fwd = P.reshape(rows // 128, 128, kb, 64).any(3).any(1)
    tm, kbw = (D + 1 + 127) // 128, (2 * B + 63) // 64
    tot = 0
    for g in range(2):
        Q = np.zeros((kbw * 64, tm * 128), dtype=bool)
        Q[:2 * B, :kb * 64] = P[g * B:g * B + 2 * B]
        tot += int(Q.reshape(kbw, 64, tm, 128).any(3).any(1).sum())
    return float(fwd.sum()) / (fwd.shape[0] * kb), tot / (2.0 * tm * kbw)


def real_images():
    """(locks, keys) of the reference's overlap_micro.zip as bool [100, 200, 200]."""
    g = np.load(os.path.join(GOLDEN, "overlap_micro_100.npz"))
    L = np.unpackbits(g["lock_bits"], axis=1)[:, :40000].reshape(-1, 200, 200).astype(bool)
    K = np.unpackbits(g["key_bits"], axis=1)[:, :40000].reshape(-1, 200, 200).astype(bool)
    return L, K


def real_batch_planes(B: int, seed: int = 0):
    """A batch of B pairs sampled with replacement from the real images, third plane = Q14 rotation of the lock by an
    angle ~ U[0, 2 pi) (8c/utils.py:329-340): bool [3B, 40000] as rows [rot; lock; key]."""
    L, K = real_images()
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(L), B)
    th = rng.uniform(0.0, 2.0 * np.pi, B)
    rot = np.stack([nn_rotate(L[i], int(round(np.cos(t) * 16384)), int(round(np.sin(t) * 16384))) for i, t in zip(idx, th)])
    return np.concatenate([rot, L[idx], K[idx]]).reshape(3 * B, -1)


def shape_stats(imgs: np.ndarray):
    """Per-image area, bounding-box height/width, centroid (y, x) of bool [n, side, side]."""
    n = len(imgs)
    area = imgs.reshape(n, -1).sum(1)
    h, w, cy, cx = np.zeros(n), np.zeros(n), np.zeros(n), np.zeros(n)
    for i, a in enumerate(imgs):
        ys, xs = np.nonzero(a)
        if len(ys) == 0:
            continue
        h[i], w[i] = ys.max() - ys.min() + 1, xs.max() - xs.min() + 1
        cy[i], cx[i] = ys.mean(), xs.mean()
    return area, h, w, cy, cx
