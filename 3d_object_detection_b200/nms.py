"""Standup-box NMS ("next" row N4) backed by ``pf_nms`` (csrc/pf_nms.cu).

Mirrors of what the reference's post-processing calls once per class and frame:

* ``nms_gpu(dets, nms_overlap_thresh, device_id=0)`` -- framework/nms.py:6-44: ``dets`` (N, 5) = x0, y0, x1, y1,
  score; returns the list of kept row indices in descending score order;
* ``nms(bboxes, scores, pre_max_size, post_max_size, iou_threshold)`` -- framework/inference.py:689-703: int64
  array of at most ``post_max_size`` kept indices, or ``None`` when nothing is kept;
* ``nms_device(boxes, scores, seg_offsets, ...)`` -- the batched form the reference lacks: all (class, frame)
  segments of a batch in three launches, device in, device out, no synchronisation.

The reference argsorts on the host, uploads, downloads an (N x N/64) bit mask and walks it in a numba loop; here the
sort, the mask and the walk run on the device and only the kept indices come back.  Same arithmetic (IoU with the
reference's ``+ 1`` box sides, float64 quotient of float32 differences as compiled numba types it, strict ``>``).
No CPU fallback.
"""
from __future__ import annotations

import numpy as np

from . import _lib

MAX_SEGMENT_BOXES = 4096


def _torch():
    import torch
    return torch


_scratch = {}


def _scratch_for(dev, n_seg, max_seg):
    torch = _torch()
    lib = _lib.load()
    need = int(lib.pf_nms_scratch_bytes(int(n_seg), int(max_seg)))
    if need == 0:
        raise ValueError(f"unsupported NMS size: {n_seg} segments of up to {max_seg} boxes (limit {MAX_SEGMENT_BOXES})")
    key = (dev.index, _lib.raw_stream(dev.index))
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty((need,), dtype=torch.uint8, device=dev)
        _scratch[key] = buf
    return buf


def nms_device(boxes, scores, seg_offsets=None, max_segment_boxes=None, iou_threshold=0.5, post_max_size=300,
               keep=None, keep_num=None):
    """boxes CUDA fp32 (N, 4), scores CUDA fp32 (N,), seg_offsets CUDA int32 (S + 1,) or None (one segment).
    Returns (keep (S, post_max_size) int32, -1 padded, indices local to the segment; keep_num (S,) int32)."""
    torch = _torch()
    if not torch.cuda.is_available():
        raise _lib.PillarFrontError("no CUDA device: the NMS has no CPU fallback")
    if boxes.dtype != torch.float32 or scores.dtype != torch.float32 or not boxes.is_cuda or not scores.is_cuda:
        raise TypeError("boxes / scores must be CUDA float32 tensors")
    boxes = boxes.contiguous()
    scores = scores.contiguous()
    if boxes.dim() != 2 or boxes.shape[1] != 4 or scores.shape[0] != boxes.shape[0]:
        raise ValueError(f"boxes must be (N, 4) and scores (N,), got {tuple(boxes.shape)} / {tuple(scores.shape)}")
    dev = boxes.device
    n_total = int(boxes.shape[0])
    n_seg = 1 if seg_offsets is None else int(seg_offsets.shape[0]) - 1
    if max_segment_boxes is None:
        if seg_offsets is not None:
            raise ValueError("max_segment_boxes is required with seg_offsets (it sizes the launch without a sync)")
        max_segment_boxes = max(n_total, 1)
    post = int(post_max_size) if post_max_size is not None else int(max_segment_boxes)
    post = max(post, 1)
    lib = _lib.load()
    if keep is None:
        keep = torch.empty((n_seg, post), dtype=torch.int32, device=dev)
    if keep_num is None:
        keep_num = torch.empty((n_seg,), dtype=torch.int32, device=dev)
    with _lib.on_device(dev):
        scratch = _scratch_for(dev, n_seg, max_segment_boxes)
        _lib.check(lib.pf_nms(boxes.data_ptr() if n_total else None, scores.data_ptr() if n_total else None,
                              seg_offsets.data_ptr() if seg_offsets is not None else None, n_seg, n_total,
                              int(max_segment_boxes), float(iou_threshold), post, scratch.data_ptr(), keep.data_ptr(),
                              keep_num.data_ptr(), _lib.raw_stream(dev.index)), "pf_nms")
    return keep, keep_num


def nms_gpu(dets, nms_overlap_thresh, device_id=0):
    """Same signature and return value (list of kept row indices) as framework/nms.py:nms_gpu."""
    torch = _torch()
    if not torch.cuda.is_available():
        raise _lib.PillarFrontError("no CUDA device: the NMS has no CPU fallback")
    dev = torch.device("cuda", int(device_id))
    if isinstance(dets, np.ndarray):
        d = torch.from_numpy(np.ascontiguousarray(dets, np.float32)).to(dev)
    else:
        d = dets.to(device=dev, dtype=torch.float32)
    n = int(d.shape[0])
    if n == 0:
        return []
    if n > MAX_SEGMENT_BOXES:
        raise ValueError(f"nms_gpu: {n} boxes > {MAX_SEGMENT_BOXES} (the reference caps at pre_max_size = 1000)")
    keep, num = nms_device(d[:, :4].contiguous(), d[:, 4].contiguous(), None, n, float(nms_overlap_thresh), n)
    k = int(num.item())
    return keep[0, :k].cpu().numpy().tolist()


def nms(bboxes, scores, pre_max_size=None, post_max_size=None, iou_threshold=0.5):
    """framework/inference.py:689-703.  numpy inputs give a numpy int64 array (or None); CUDA tensors give a CUDA int64
    tensor (or None)."""
    torch = _torch()
    is_t = hasattr(bboxes, "is_cuda")
    n = int(bboxes.shape[0])
    if n == 0:
        return None
    if is_t:
        b = bboxes.to(dtype=torch.float32)
        s = scores.to(device=b.device, dtype=torch.float32)
    else:
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
        if dev is None:
            raise _lib.PillarFrontError("no CUDA device: the NMS has no CPU fallback")
        b = torch.from_numpy(np.ascontiguousarray(bboxes, np.float32)).to(dev)
        s = torch.from_numpy(np.ascontiguousarray(scores, np.float32)).to(dev)
    if n > MAX_SEGMENT_BOXES:
        raise ValueError(f"nms: {n} boxes > {MAX_SEGMENT_BOXES}; apply the top-k (pre_max_size) first, as the reference does")
    post = n if post_max_size is None else min(int(post_max_size), n)
    keep, num = nms_device(b, s, None, n, float(iou_threshold), max(post, 1))
    k = int(num.item())
    if k == 0 or post == 0:
        return None
    out = keep[0, :k].to(torch.int64)
    return out if is_t else out.cpu().numpy()
