"""Seeded synthetic LiDAR frames (S1 clustered, S2 ring, S3 uniform box, S4 edge set).

All generators return C-contiguous float32 ``(N, 4)`` arrays ``[x, y, z, intensity]`` built from
``numpy.random.default_rng(seed)``; there is no dataset on the box (BASELINE.json: synthetic frames).
"""
from __future__ import annotations

import numpy as np

from .config import GridGeometry

# occupied-cell targets for S1 per grid key
S1_CELLS = {"20cm": 12000, "10cm": 15000, "nuscene": 9000, "16cm": 12000}


def frame_seed(config_id: int, frame_id: int) -> int:
    return 1000 * int(config_id) + int(frame_id)


def s1_clustered(geom: GridGeometry, n: int, seed: int, n_cells: int = 12000) -> np.ndarray:
    """S1: ``n_cells`` distinct xy cells, heavy-tailed weights ``k^-0.8``; ~8 % of points fall outside in z."""
    rng = np.random.default_rng(seed)
    nx, ny = int(geom.grid_size[0]), int(geom.grid_size[1])
    n_cells = min(n_cells, nx * ny)
    cells = rng.choice(nx * ny, size=n_cells, replace=False)
    w = np.arange(1, n_cells + 1, dtype=np.float64) ** -0.8
    w /= w.sum()
    pick = cells[rng.choice(n_cells, size=n, p=w)]
    cx = (pick // ny).astype(np.float64)
    cy = (pick % ny).astype(np.float64)
    vs = geom.voxel_size.astype(np.float64)
    off = geom.offset.astype(np.float64)
    pts = np.empty((n, 4), dtype=np.float32)
    pts[:, 0] = off[0] + (cx + rng.uniform(0.05, 0.95, n)) * vs[0]
    pts[:, 1] = off[1] + (cy + rng.uniform(0.05, 0.95, n)) * vs[1]
    zmin, zmax = float(geom.detection_range[2]), float(geom.detection_range[5])
    pts[:, 2] = rng.uniform(zmin - 0.5, zmax + 0.5, n)
    pts[:, 3] = rng.uniform(0.0, 1.0, n)
    return pts


def s2_ring(geom: GridGeometry, n: int, seed: int, beams: int = 64) -> np.ndarray:
    """S2: spinning 64-beam LiDAR over flat ground, azimuth-major point order (exercises the max_voxels break)."""
    rng = np.random.default_rng(seed)
    steps = max(1, n // beams)
    n_eff = steps * beams
    elev = np.deg2rad(np.linspace(-24.8, 2.0, beams))
    az = np.linspace(0.0, 2.0 * np.pi, steps, endpoint=False)
    azg, elg = np.meshgrid(az, elev, indexing="ij")            # azimuth-major
    azg, elg = azg.ravel(), elg.ravel()
    sensor_h = 1.9
    with np.errstate(divide="ignore", invalid="ignore"):
        ground = np.where(elg < 0, sensor_h / np.maximum(np.sin(-elg), 1e-9), np.inf)
    obstacle = 2.0 + rng.gamma(2.0, 15.0, n_eff)
    rng_m = np.minimum(ground, obstacle)
    rng_m *= 1.0 + 0.002 * rng.standard_normal(n_eff)
    pts = np.empty((n_eff, 4), dtype=np.float32)
    pts[:, 0] = rng_m * np.cos(elg) * np.cos(azg)
    pts[:, 1] = rng_m * np.cos(elg) * np.sin(azg)
    pts[:, 2] = sensor_h - 1.9 + rng_m * np.sin(elg)
    pts[:, 3] = rng.uniform(0.0, 1.0, n_eff)
    return pts


def s3_uniform(geom: GridGeometry, n: int, seed: int) -> np.ndarray:
    """S3: uniform over 1.1x the xy range (≈17 % outside); nearly every inside point opens a new cell."""
    rng = np.random.default_rng(seed)
    lo = geom.detection_range[:3].astype(np.float64)
    hi = geom.detection_range[3:].astype(np.float64)
    mid, half = (lo + hi) / 2, (hi - lo) / 2
    pts = np.empty((n, 4), dtype=np.float32)
    pts[:, 0] = rng.uniform(mid[0] - 1.1 * half[0], mid[0] + 1.1 * half[0], n)
    pts[:, 1] = rng.uniform(mid[1] - 1.1 * half[1], mid[1] + 1.1 * half[1], n)
    pts[:, 2] = rng.uniform(lo[2], hi[2], n)
    pts[:, 3] = rng.uniform(0.0, 1.0, n)
    return pts


def edge_points(geom: GridGeometry) -> np.ndarray:
    """S4 micro-cases: exact boundaries, signed zeros, non-finite and out-of-int32 values."""
    lo = geom.detection_range[:3]
    hi = geom.detection_range[3:]
    f = np.float32
    mid = ((lo + hi) / 2).astype(np.float32)
    rows = [
        [lo[0], lo[1], lo[2], 0.1],                                # lower corner: inside, cell (0,0,0)
        [hi[0], hi[1], hi[2], 0.2],                                # upper corner: outside (exclusive)
        [np.nextafter(hi[0], f(-np.inf)), np.nextafter(hi[1], f(-np.inf)), np.nextafter(hi[2], f(-np.inf)), 0.3],
        [np.nextafter(lo[0], f(-np.inf)), mid[1], mid[2], 0.4],    # just below the lower bound
        [f(-0.0), f(0.0), mid[2], 0.5],
        [np.nan, mid[1], mid[2], 0.6],
        [mid[0], np.inf, mid[2], 0.7],
        [mid[0], mid[1], -np.inf, 0.8],
        [3e9, mid[1], mid[2], 0.9],
        [-3e9, mid[1], mid[2], 1.0],
        [f(2.0 ** 32) * geom.voxel_size[0] + lo[0], mid[1], mid[2], 0.15],   # floor(...) = 2^32: not cell 0
        [f(3.4e38), f(-3.4e38), mid[2], 0.25],
        [mid[0], mid[1], mid[2], np.nan],                          # NaN intensity is payload, not geometry
        [mid[0], mid[1], mid[2], 0.35],                            # same cell as the row above
    ]
    return np.asarray(rows, dtype=np.float32)


def s4_edges(geom: GridGeometry, n: int, seed: int, n_cells: int = 12000) -> np.ndarray:
    """S4: an S1 frame with the edge set spliced in at the front, the middle and the end."""
    base = s1_clustered(geom, max(n, 8), seed, n_cells)
    e = edge_points(geom)
    h = base.shape[0] // 2
    return np.ascontiguousarray(np.concatenate([e, base[:h], e[::-1], base[h:], e], axis=0))


def make_frame(kind: str, geom: GridGeometry, n: int, seed: int, n_cells: int = 12000) -> np.ndarray:
    kind = kind.upper()
    if kind == "S1":
        return s1_clustered(geom, n, seed, n_cells)
    if kind == "S2":
        return s2_ring(geom, n, seed)
    if kind == "S3":
        return s3_uniform(geom, n, seed)
    if kind == "S4":
        return s4_edges(geom, n, seed, n_cells)
    raise ValueError(f"unknown synthetic frame kind {kind!r}")


# sizes (l, w, h) and yaw angles of the anchor set the reference hard-codes (framework/anchor_assigner.py:222-246)
ANCHOR_CLASSES = (
    ((4.6, 2.10, 1.8), (0.0, 1.5707963267948966)),
    ((7.5, 2.6, 2.9), (0.0, 1.5707963267948966)),
    ((12.6, 2.9, 3.8), (0.0, 1.5707963267948966)),
    ((0.96874749, 0.9645992, 1.81212425), (0.0,)),
    ((2.02032733, 0.98075615, 1.72027404), (0.0, 1.5707963267948966)),
)


def anchors_bv(geom: GridGeometry, feature_map=(400, 400), classes=ANCHOR_CLASSES) -> np.ndarray:
    """Synthetic bird's-eye anchor boxes (A, 4) fp32 = (xmin, ymin, xmax, ymax): one box per feature-map cell, size
    and yaw, centred on the cell, length and width swapped for the 90-degree yaw.  Same population (9 boxes per
    feature-map cell = 1.44 M at 400 x 400) as the reference's AnchorAssigner builds at start-up."""
    fx, fy = int(feature_map[0]), int(feature_map[1])
    diff = np.asarray(geom.range_diff, np.float32)
    off = np.asarray(geom.offset, np.float32)
    sx, sy = diff[0] / np.float32(fx), diff[1] / np.float32(fy)
    cx = np.arange(fx, dtype=np.float32) * sx + (off[0] + sx / 2)
    cy = np.arange(fy, dtype=np.float32) * sy + (off[1] + sy / 2)
    gx, gy = np.meshgrid(cx, cy, indexing="ij")
    gx, gy = gx.reshape(-1), gy.reshape(-1)
    out = []
    for (l, w, _h), yaws in classes:
        for yaw in yaws:
            dx, dy = (np.float32(l), np.float32(w)) if abs(yaw) < 0.7 else (np.float32(w), np.float32(l))
            out.append(np.stack([gx - dx / 2, gy - dy / 2, gx + dx / 2, gy + dy / 2], axis=1))
    return np.ascontiguousarray(np.concatenate(out, axis=0), np.float32)


def nms_dets(n: int, seed: int) -> np.ndarray:
    """(n, 5) fp32 detections = standup box (x0, y0, x1, y1) + distinct score: car-sized boxes clustered around a few
    centres, so that suppression chains occur (input of the NMS row N4)."""
    rng = np.random.default_rng(seed)
    k = max(1, n // 6)
    centres = rng.uniform(-40, 40, (k, 2))
    c = centres[rng.integers(0, k, n)] + rng.normal(0, 1.2, (n, 2))
    wh = np.abs(rng.normal([4.5, 2.0], [0.8, 0.4], (n, 2))) + 0.2
    swap = rng.random(n) < 0.5
    wh[swap] = wh[swap][:, ::-1]
    boxes = np.concatenate([c - wh / 2, c + wh / 2], 1)
    scores = rng.permutation(n).astype(np.float64) / n * 0.9 + 0.05 + rng.uniform(0, 1e-3, n)   # distinct scores
    return np.concatenate([boxes, scores[:, None]], 1).astype(np.float32)


def anchors_3d(geom: GridGeometry, feature_map=(400, 400), classes=ANCHOR_CLASSES):
    """Synthetic 7-number anchors (A, 7) fp32 = (x, y, z, l, w, h, yaw), one per feature-map cell, size and yaw, laid
    out like AnchorAssigner builds them (framework/anchor_assigner.py:252-289, generate :301-320): all cells of one
    (size, yaw) pair contiguously, x slow.  Also returns the class ranges [[start, end), ...] (class_masks) of the
    reference's three classes: the first three sizes are "vehicle", then "pedestrian", then "cyclist"."""
    fx, fy = int(feature_map[0]), int(feature_map[1])
    diff = np.asarray(geom.range_diff, np.float32)
    off = np.asarray(geom.offset, np.float32)
    sx, sy = diff[0] / np.float32(fx), diff[1] / np.float32(fy)
    cx = np.arange(fx, dtype=np.float32) * sx + (off[0] + sx / 2)
    cy = np.arange(fy, dtype=np.float32) * sy + (off[1] + sy / 2)
    gx, gy = np.meshgrid(cx, cy, indexing="ij")
    gx, gy = gx.reshape(-1), gy.reshape(-1)
    out, per_size = [], []
    for (l, w, h), yaws in classes:
        n0 = sum(len(o) for o in out)
        for yaw in yaws:
            a = np.empty((gx.size, 7), np.float32)
            a[:, 0], a[:, 1], a[:, 2] = gx, gy, np.float32(h / 2)
            a[:, 3], a[:, 4], a[:, 5], a[:, 6] = l, w, h, yaw
            out.append(a)
        per_size.append((n0, sum(len(o) for o in out)))
    anchors = np.ascontiguousarray(np.concatenate(out, axis=0))
    if len(per_size) == 5:          # the reference's set: three vehicle sizes, pedestrian, cyclist
        ranges = [[per_size[0][0], per_size[2][1]], list(per_size[3]), list(per_size[4])]
    else:
        ranges = [list(r) for r in per_size]
    return anchors, ranges


def post_preds(n_frames: int, n_anchors: int, seed: int, mean_logit: float = -6.0, mask_frac: float = 0.7):
    """Synthetic network-head outputs for the post-processing row: class logits (B, A, 1) ~ N(mean_logit, 2) (about 6 %
    above the 0.05 score threshold at -6), box encodings (B, A, 7) ~ N(0, 0.25) with yaw residuals U(-pi, pi), direction
    logits (B, A, 2), and an anchor mask (B, A) bool."""
    rng = np.random.default_rng(seed)
    cls = rng.normal(mean_logit, 2.0, (n_frames, n_anchors, 1)).astype(np.float32)
    box = rng.normal(0.0, 0.25, (n_frames, n_anchors, 7)).astype(np.float32)
    box[..., 6] = rng.uniform(-np.pi, np.pi, (n_frames, n_anchors)).astype(np.float32)
    dirp = rng.normal(0.0, 1.0, (n_frames, n_anchors, 2)).astype(np.float32)
    mask = rng.random((n_frames, n_anchors)) < mask_frac
    return cls, box, dirp, mask


def post_threshold_band(cls: np.ndarray, score_threshold: float = 0.05, half_width: float = 0.01, seed: int = 0) -> np.ndarray:
    """Overwrite the class logits (in place) with values spread over [L - half_width, L + half_width], L = the logit of the
    score threshold: the region where the score mask is decided by the computed sigmoid itself.  Logits whose fp64 sigmoid
    lies within 2e-7 of the threshold (where two correct fp32 implementations may disagree) are moved up by half_width / 2."""
    rng = np.random.default_rng(seed)
    L = np.log(score_threshold / (1.0 - score_threshold))
    x = rng.uniform(L - half_width, L + half_width, cls.shape).astype(np.float32)
    s = 1.0 / (1.0 + np.exp(-x.astype(np.float64)))
    x[np.abs(s - np.float64(np.float32(score_threshold))) < 2e-7] += np.float32(half_width / 2)
    cls[...] = x
    return cls
