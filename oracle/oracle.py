"""ORACLE -- TEST INFRASTRUCTURE ONLY.  ctypes front of oracle/pillar_oracle.c plus numpy/torch mirrors.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module, and only as the checker or the timed CPU baseline.  The product package
(3d_object_detection_b200/) never imports it and fails loudly when its CUDA library is missing.

Parity status: PINNED against outputs of the reference itself (tests/golden/, made by
oracle/make_golden.py from /root/reference; checked by tests/test_oracle.py).
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from . import build as _build

_LIB = None

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def lib():
    global _LIB
    if _LIB is None:
        path = _build.LIB if _build.LIB.exists() else _build.build()
        L = C.CDLL(str(path))
        L.oracle_points_to_voxels.restype = C.c_int32
        L.oracle_points_to_voxels.argtypes = [_f32p, C.c_int64, C.c_int32, _f32p, _i32p, _i32p, _i32p,
                                              _f32p, _f32p, _i32p, C.c_int32, C.c_int32]
        L.oracle_generate.restype = C.c_int32
        L.oracle_generate.argtypes = [_f32p, C.c_int64, C.c_int32, _f32p, _i32p, _i32p,
                                      _f32p, _f32p, _i32p, C.c_int32, C.c_int32]
        L.oracle_decorate.restype = None
        L.oracle_decorate.argtypes = [_f32p, _i32p, _i32p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_float, C.c_float, C.c_float, C.c_float, _f32p]
        L.oracle_scatter.restype = None
        L.oracle_scatter.argtypes = [_f32p, _i32p, C.c_int32, C.c_int64, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_int32, _f32p]
        L.oracle_scatter_cmajor.restype = None
        L.oracle_scatter_cmajor.argtypes = [_f32p, _i32p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                            C.c_int32, _f32p]
        L.oracle_frontend_batch.restype = C.c_int64
        L.oracle_frontend_batch.argtypes = [_f32p, _i64p, C.c_int32, C.c_int32, _f32p, _f32p, _i32p,
                                            C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_float,
                                            C.c_float, _f32p, C.c_int32, _f32p, _i32p, _i32p, _i32p,
                                            _f32p, _f32p, C.c_int32]
        L.oracle_nms_iou.restype = None
        L.oracle_nms_iou.argtypes = [_f32p, C.c_int32, np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")]
        _LIB = L
    return _LIB


def _geom_arrays(geom):
    return (np.ascontiguousarray(geom.voxel_size, np.float32), np.ascontiguousarray(geom.offset, np.float32),
            np.ascontiguousarray(geom.grid_size, np.int32))


def generate(geom, points, max_voxels, max_points):
    """VoxelGenerator.generate (framework/voxel_generator.py:28-40): returns (voxels, coors, num) sliced to V."""
    points = np.ascontiguousarray(points, np.float32)
    n, f = points.shape
    vs, off, gs = _geom_arrays(geom)
    voxels = np.empty((max_voxels, max_points, f), np.float32)
    num = np.empty((max_voxels,), np.int32)
    coors = np.empty((max_voxels, 3), np.int32)
    v = lib().oracle_generate(points.reshape(-1) if n else np.zeros(1, np.float32), n, f,
                              voxels.reshape(-1), num, coors.reshape(-1), vs, off, gs, max_voxels, max_points)
    if v < 0:
        raise MemoryError("oracle_generate: scratch grid allocation failed")
    return voxels[:v], coors[:v], num[:v]


def decorate_consts(voxel_size_cfg, offset):
    """vx, vy, x_offset, y_offset exactly as PointNet.__init__ forms them (pointpillars8_shared.py:16-19):
    vx, vy come from the *config list* (Python floats), the offsets from numpy fp32 (F13)."""
    vx = voxel_size_cfg[0]
    vy = voxel_size_cfg[1]
    x_off = vx / 2 + offset[0]
    y_off = vy / 2 + offset[1]
    return float(np.float32(vx)), float(np.float32(vy)), float(np.float32(x_off)), float(np.float32(y_off))


def decorate(voxels, num, coors, vx, vy, x_off, y_off):
    """C oracle of pointpillars8_shared.py:32-54 -> (V,P,F+5) fp32."""
    voxels = np.ascontiguousarray(voxels, np.float32)
    num = np.ascontiguousarray(num, np.int32)
    coors = np.ascontiguousarray(coors, np.int32)
    v, p, f = voxels.shape
    out = np.empty((v, p, f + 5), np.float32)
    if v:
        lib().oracle_decorate(voxels.reshape(-1), num, coors.reshape(-1), coors.shape[1], v, p, f,
                              vx, vy, x_off, y_off, out.reshape(-1))
    return out


def decorate_numpy(voxels, num, coors, vx, vy, x_off, y_off):
    """numpy fp32 mirror of the same op sequence as the torch code (sum over axis 1, divide, subtract,
    multiply-then-add centre, concatenate, mask multiply) -- a second opinion on the C oracle."""
    f32 = np.float32
    voxels = np.asarray(voxels, f32)
    mean = voxels[:, :, :3].sum(axis=1, keepdims=True, dtype=f32) / num.astype(f32).reshape(-1, 1, 1)
    f_cluster = voxels[:, :, :3] - mean
    f_center = np.zeros_like(voxels[:, :, :2])
    f_center[:, :, 0] = voxels[:, :, 0] - (coors[:, 0].astype(f32)[:, None] * f32(vx) + f32(x_off))
    f_center[:, :, 1] = voxels[:, :, 1] - (coors[:, 1].astype(f32)[:, None] * f32(vy) + f32(y_off))
    feats = np.concatenate([voxels, f_cluster, f_center], axis=-1)
    mask = (num.reshape(-1, 1) > np.arange(voxels.shape[1], dtype=np.int32).reshape(1, -1))
    feats = feats * mask[..., None].astype(f32)
    return feats


def decorate_f64(voxels, num, coors, vx, vy, x_off, y_off):
    """fp64 evaluation of the decoration: the yardstick both fp32 versions are held to."""
    v64 = np.asarray(voxels, np.float64)
    mean = v64[:, :, :3].sum(axis=1, keepdims=True) / num.astype(np.float64).reshape(-1, 1, 1)
    fc = v64[:, :, :3] - mean
    cx = coors[:, 0].astype(np.float64)[:, None] * float(np.float32(vx)) + float(np.float32(x_off))
    cy = coors[:, 1].astype(np.float64)[:, None] * float(np.float32(vy)) + float(np.float32(y_off))
    ce = np.stack([v64[:, :, 0] - cx, v64[:, :, 1] - cy], axis=-1)
    feats = np.concatenate([v64, fc, ce], axis=-1)
    mask = (num.reshape(-1, 1) > np.arange(v64.shape[1]).reshape(1, -1))
    return feats * mask[..., None]


def decorate_tolerance(voxels):
    """Per-pillar absolute tolerance of the decoration parity check: 1e-6 relative to the
    operand magnitude, i.e. 1e-6 * max(1, max |xyz| in the pillar), shape (V,1,1)."""
    voxels = np.asarray(voxels)
    if len(voxels) == 0:
        return np.zeros((0, 1, 1), np.float32)
    mag = np.abs(voxels[:, :, :3]).reshape(len(voxels), -1).max(axis=1)
    return (1e-6 * np.maximum(1.0, mag)).reshape(-1, 1, 1)


def scatter(feats, coors, nx, ny, batch=1):
    """PointPillarsScatter.forward (pointpillars8_shared.py:76-111), row-major (V,C) features -> (B,C,nx,ny)."""
    feats = np.ascontiguousarray(feats, np.float32)
    coors = np.ascontiguousarray(coors, np.int32)
    v, c = feats.shape
    canvas = np.empty((batch, c, nx, ny), np.float32)
    lib().oracle_scatter(feats.reshape(-1) if v else np.zeros(1, np.float32),
                         coors.reshape(-1) if v else np.zeros(4, np.int32),
                         coors.shape[1] if v else 3, v, c, nx, ny, batch, canvas.reshape(-1))
    return canvas


def scatter_cmajor(feats_cv, coors, nx, ny):
    """Scatter.run (pointpillars8_export.py:44-51), channel-major (C,V) features -> (1,C,nx,ny)."""
    feats_cv = np.ascontiguousarray(feats_cv, np.float32)
    coors = np.ascontiguousarray(coors, np.int32)
    c, v = feats_cv.shape
    canvas = np.empty((1, c, nx, ny), np.float32)
    lib().oracle_scatter_cmajor(feats_cv.reshape(-1) if v else np.zeros(1, np.float32),
                                coors.reshape(-1) if v else np.zeros(4, np.int32),
                                coors.shape[1] if v else 3, v, c, nx, ny, canvas.reshape(-1))
    return canvas


def scatter_backward(grad_canvas, coors, batch=1):
    """Gradient of the scatter w.r.t. voxel_features (autograd of index_put): a gather."""
    g = np.asarray(grad_canvas, np.float32)
    b, c, nx, ny = g.shape
    coors = np.asarray(coors)
    bi = coors[:, 3] if coors.shape[1] >= 4 else np.zeros(len(coors), np.int64)
    return np.ascontiguousarray(g[bi, :, coors[:, 0], coors[:, 1]])


class FrontendBatch:
    """Reusable buffers + call of oracle_frontend_batch (bench.py CPU baseline body)."""

    def __init__(self, geom, max_voxels, max_points, n_feat, n_frames, n_ch=64):
        self.geom, self.mv, self.mp, self.f, self.b, self.c = geom, max_voxels, max_points, n_feat, n_frames, n_ch
        nx, ny = int(geom.grid_size[0]), int(geom.grid_size[1])
        self.voxels = np.empty((n_frames, max_voxels, max_points, n_feat), np.float32)
        self.num = np.empty((n_frames, max_voxels), np.int32)
        self.coors = np.empty((n_frames, max_voxels, 3), np.int32)
        self.voxel_num = np.zeros((n_frames,), np.int32)
        self.features = np.empty((n_frames, max_voxels, max_points, n_feat + 5), np.float32)
        self.canvas = np.empty((n_frames, n_ch, nx, ny), np.float32)

    def run(self, points_cat, frame_offsets, consts, pillar_feats, n_threads):
        vs, off, gs = _geom_arrays(self.geom)
        vx, vy, xo, yo = consts
        tot = lib().oracle_frontend_batch(
            np.ascontiguousarray(points_cat, np.float32).reshape(-1),
            np.ascontiguousarray(frame_offsets, np.int64), self.b, self.f, vs, off, gs, self.mv, self.mp,
            vx, vy, xo, yo, np.ascontiguousarray(pillar_feats, np.float32).reshape(-1), self.c,
            self.voxels.reshape(-1), self.num.reshape(-1), self.coors.reshape(-1), self.voxel_num,
            self.features.reshape(-1), self.canvas.reshape(-1), int(n_threads))
        if tot < 0:
            raise MemoryError("oracle_frontend_batch: allocation failed")
        return int(tot)


def pfn(features, weight, gamma, beta, mean, var, eps=1e-5):
    """fp64 restatement of the learned tail of PointNet.forward in eval mode (pointpillars8_shared.py:24-28,57-60):
    Conv1d(9->64, k=1, no bias) + BatchNorm1d(running stats) + ReLU + max over the P slots -> (V, 64)."""
    f = np.asarray(features, np.float64)                      # (V, P, 9)
    w = np.asarray(weight, np.float64).reshape(-1, f.shape[2])  # (64, 9)
    y = f @ w.T                                               # (V, P, 64)
    y = (y - np.asarray(mean, np.float64)) / np.sqrt(np.asarray(var, np.float64) + eps) * np.asarray(gamma, np.float64) \
        + np.asarray(beta, np.float64)
    return np.maximum(y, 0.0).max(axis=1)


def anchor_coor(anchors_bv, stride, offset, grid_size):
    """get_anchor_coor, framework/box_np_ops.py:288-304: floor((edge - offset) / stride) in the arrays' own
    precision, stored as int32, minima clipped at 0 and maxima at grid - 1 (and nothing else)."""
    bv = np.asarray(anchors_bv)
    out = np.zeros(bv.shape, dtype=np.int32)
    for k, axis in enumerate((0, 1, 0, 1)):
        out[:, k] = np.floor((bv[:, k] - offset[axis]) / stride[axis]).astype(np.int32)
    out[:, 0] = np.maximum(out[:, 0], 0)
    out[:, 1] = np.maximum(out[:, 1], 0)
    out[:, 2] = np.minimum(out[:, 2], grid_size[0] - 1)
    out[:, 3] = np.minimum(out[:, 3], grid_size[1] - 1)
    return out


def anchors_mask(coors, shape, anchor_coors):
    """AnchorAssigner.create_mask, framework/anchor_assigner.py:322-335, in the arithmetic both of its branches
    share: fp32 pillar counts per (x, y) (sparse_sum_for_anchors_mask, box_np_ops.py:160-165 = init_map_gpu
    :168-172), running sums along axis 0 then axis 1 (cumx_gpu / cumy_gpu :175-188), and the four-corner test of
    get_anchors_mask_gpu (:214-223) / fused_get_anchors_area (:261-285) on the anchor rectangles."""
    dense = np.zeros(tuple(int(s) for s in shape), dtype=np.float32)
    coors = np.asarray(coors)
    for i in range(coors.shape[0]):
        dense[coors[i, 0], coors[i, 1]] += 1
    dense = dense.cumsum(0)
    dense = dense.cumsum(1)
    minx, miny, maxx, maxy = (np.asarray(anchor_coors)[:, k] for k in range(4))
    area = dense[maxx, maxy] - dense[maxx, miny] - dense[minx, maxy] + dense[minx, miny]
    return area > 0


def nms_iou_matrix(boxes):
    """iou_device(a = row box, b = column box), framework/nms.py:102-113, for all pairs, in the arithmetic compiled numba
    gives it (oracle_nms_iou in pillar_oracle.c; PTX evidence tests/golden/nms_iou_ptx_ops.json): the box edges are float32,
    `right - left` is a float32 subtraction, `+ 1` (an int64 literal) promotes everything after it to float64 (numba typing:
    float32 + int64 -> float64), and Sa + Sb is one fma."""
    b = np.ascontiguousarray(np.asarray(boxes, np.float32)[:, :4])
    n = b.shape[0]
    out = np.empty((n, n), np.float64)
    if n:
        lib().oracle_nms_iou(b.reshape(-1), n, out.reshape(-1))
    return out


def nms_gpu(dets, nms_overlap_thresh):
    """nms_gpu + nms_kernel + nms_postprocess, framework/nms.py:6-44, 85-147: boxes sorted by descending score, a
    box is dropped when an earlier KEPT box overlaps it by more than the threshold (strict >).  Returns indices into
    `dets` in keep order.  Ties in the score are ordered like a stable ascending argsort read backwards."""
    dets = np.asarray(dets, np.float32)
    n = dets.shape[0]
    order = np.argsort(dets[:, 4], kind="stable")[::-1].astype(np.int32)
    iou = nms_iou_matrix(dets[order, :4])
    over = iou > np.float64(np.float32(nms_overlap_thresh))
    removed = np.zeros(n, bool)
    keep = []
    for i in range(n):
        if not removed[i]:
            keep.append(i)
            removed[i + 1:] |= over[i, i + 1:]
    return [int(order[k]) for k in keep]


def nms(bboxes, scores, pre_max_size=None, post_max_size=None, iou_threshold=0.5):
    """nms, framework/inference.py:689-703."""
    dets = np.concatenate([bboxes, scores[:, np.newaxis]], axis=1)
    if len(dets) == 0:
        keep = np.array([], dtype=np.int64)
    else:
        keep = np.array(nms_gpu(dets, iou_threshold), dtype=np.int64)[:post_max_size]
    return None if keep.shape[0] == 0 else keep


# ---- detection post-processing (row N4 around the NMS): Inference.infer_gpu, framework/inference.py:26-138 -----------

def sigmoid32(x):
    """torch.sigmoid in fp32, written as 1 / (1 + exp(-x)) (framework/inference.py:50)."""
    x = np.asarray(x, np.float32)
    with np.errstate(over="ignore"):
        return (np.float32(1.0) / (np.float32(1.0) + np.exp(-x))).astype(np.float32)


def box_decode(box_encodings, anchors):
    """box_decode, framework/box_np_ops.py:406-424 (the live definition; the one at :386 is inside a string)."""
    xa, ya, za, la, wa, ha, ra = np.split(anchors, 7, axis=-1)
    xt, yt, zt, lt, wt, ht, rt = np.split(box_encodings, 7, axis=-1)
    za = za + ha / 2
    diagonal = np.sqrt(la ** 2 + wa ** 2)
    xg = xt * diagonal + xa
    yg = yt * diagonal + ya
    zg = zt * ha + za
    lg = np.exp(lt) * la
    wg = np.exp(wt) * wa
    hg = np.exp(ht) * ha
    rg = rt + ra
    zg = zg - hg / 2
    return np.concatenate([xg, yg, zg, lg, wg, hg, rg], axis=-1)


def standup_boxes(boxes_for_nms):
    """center_to_corner_box2d (framework/box_np_ops.py:81-99: corners_nd :122-153 with origin 0.5, rotation_2d :64-78,
    + centres) followed by corner_to_standup_nd (:899-904), on rows (x, y, l, w, r)."""
    b = np.asarray(boxes_for_nms)
    centers, dims, angles = b[:, :2], b[:, 2:4], b[:, 4]
    corners_norm = np.array([[0, 0], [0, 1], [1, 1], [1, 0]], dtype=dims.dtype) - np.array(0.5, dtype=dims.dtype)
    corners = dims.reshape([-1, 1, 2]) * corners_norm.reshape([1, 4, 2])
    rot_sin, rot_cos = np.sin(angles), np.cos(angles)
    rot_mat_T = np.stack([[rot_cos, rot_sin], [-rot_sin, rot_cos]])
    corners = np.einsum('aij,jka->aik', corners, rot_mat_T)
    corners += centers.reshape([-1, 1, 2])
    return np.concatenate([np.min(corners, axis=1), np.max(corners, axis=1)], -1)


def limit_period(val, offset=0.5, period=np.pi):
    """framework/box_np_ops.py:102-103."""
    return val - np.floor(val / period + offset) * period


def post_select(cls_logits, a_mask, score_threshold, pre_max):
    """framework/inference.py:44-69 for one class range of one frame: anchor mask, sigmoid, max over the class channels,
    score mask, top-k.  Returns the local anchor indices of the selected boxes, best first, and their scores.
    The reference orders by score and leaves ties open; the order stated here is (max logit descending, anchor index
    ascending), a refinement of it that needs no exp()."""
    logits = np.asarray(cls_logits, np.float32).reshape(len(cls_logits), -1)
    with np.errstate(invalid="ignore"):
        top_logit = logits.max(axis=1)                 # NaN propagates, and a NaN score fails the threshold test
    score = sigmoid32(top_logit)
    ok = score >= np.float32(score_threshold)
    if a_mask is not None:
        ok &= np.asarray(a_mask, bool)
    idx = np.nonzero(ok)[0]
    # total order on the logits: the usual monotone float -> uint32 map (so -0.0 sorts just below +0.0)
    u = np.ascontiguousarray(top_logit[idx]).view(np.uint32).astype(np.int64)
    key = np.where(u & 0x80000000, ~u & 0xFFFFFFFF, u | 0x80000000)
    order = np.argsort(-key, kind="stable")
    sel = idx[order][:pre_max]
    return sel.astype(np.int64), score[sel]


def post_finish(boxes, top_scores, dir_labels, selected, center_limit):
    """framework/inference.py:97-114 for one class: NMS mask, direction flip, range mask, limit_period.  Returns the
    surviving boxes and scores and the range mask over `selected`."""
    boxes = np.array(boxes, np.float32)[selected]
    scores = np.asarray(top_scores)[selected]
    dir_mask = boxes[..., -1] > 0
    opp = dir_mask ^ np.asarray(dir_labels).astype(bool)[selected]
    boxes[..., -1] += np.where(opp, np.pi, 0.0)
    limit = list(center_limit)
    min_mask = np.any(boxes[:, :3] > limit[:3], axis=1)
    max_mask = np.any(boxes[:, 3:6] < limit[3:], axis=1)
    range_mask = min_mask & max_mask
    boxes = boxes[range_mask]
    boxes[..., -1] = limit_period(boxes[..., -1], period=2 * np.pi)
    return boxes, scores[range_mask], range_mask


def postprocess_frame(cls_preds, box_preds, dir_preds, anchors_mask, anchors, class_ranges, center_limit,
                      score_threshold=0.05, pre_max=1000, post_max=300, iou_threshold=0.1, detail=None):
    """The per-class loop of Inference.infer_gpu (framework/inference.py:35-121) for one frame.  Returns
    (boxes (D, 7) fp32, scores (D,) fp32, class index (D,) int32, anchor index (D,) int64) in the order of the
    reference's concatenated lists.  `detail`, if a list, receives one dict per class with the intermediate arrays."""
    anchors = np.asarray(anchors, np.float32)
    out_b, out_s, out_c, out_a = [], [], [], []
    for c, (lo, hi) in enumerate(class_ranges):
        a_mask = None if anchors_mask is None else np.asarray(anchors_mask)[lo:hi]
        sel, top_scores = post_select(np.asarray(cls_preds)[lo:hi], a_mask, score_threshold, pre_max)
        info = dict(sel=sel, scores=top_scores)
        if detail is not None:
            detail.append(info)
        if len(sel) == 0:
            continue
        enc = np.asarray(box_preds, np.float32)[lo:hi][sel]
        d = np.asarray(dir_preds, np.float32)[lo:hi][sel]
        dir_labels = (d[:, 1] > d[:, 0])                                  # torch.max(dim=-1)[1]: first maximum on a tie
        boxes = box_decode(enc, anchors[lo:hi][sel])
        standup = standup_boxes(boxes[:, [0, 1, 3, 4, 6]])
        # the boxes are already in their final order; strictly decreasing stand-in scores make nms_gpu's argsort keep it
        rank = np.arange(len(sel), 0, -1).astype(np.float32)
        selected = nms(standup, rank, pre_max, post_max, iou_threshold)
        info.update(boxes=boxes.copy(), standup=standup, dir=dir_labels.astype(np.int32),
                    keep=np.zeros(0, np.int64) if selected is None else selected)
        if selected is None:
            continue
        boxes, scores_kept, range_mask = post_finish(boxes, top_scores, dir_labels, selected, center_limit)
        out_b.append(boxes)
        out_s.append(scores_kept)
        out_c.append(np.full(len(boxes), c, np.int32))
        out_a.append(lo + sel[selected][range_mask])
    if not out_b:
        return (np.zeros((0, 7), np.float32), np.zeros((0,), np.float32), np.zeros((0,), np.int32), np.zeros((0,), np.int64))
    return (np.concatenate(out_b).astype(np.float32), np.concatenate(out_s).astype(np.float32), np.concatenate(out_c),
            np.concatenate(out_a).astype(np.int64))


def post_min_iou_gap(standup, iou_threshold):
    """Smallest |IoU - threshold| over all pairs of standup boxes (how far a case is from a float-sensitive decision)."""
    if len(standup) < 2:
        return np.inf
    iou = nms_iou_matrix(standup)
    iu = np.triu_indices(len(standup), 1)
    return float(np.abs(iou[iu] - np.float64(np.float32(iou_threshold))).min())
