"""Detection post-processing (row N4, the part around the NMS) backed by ``pf_postprocess`` (csrc/pf_post.cu).

Mirror of the reference's ``framework/inference.py``:

* ``Inference(config, anchor_assigner)`` -- same constructor (:10-24: reads ``config['device']``, ``config['center_limit']``,
  ``config['detect_class']``, ``anchor_assigner.anchors`` and ``.class_masks``), same attributes
  (``_nms_pre_max_size`` 1000, ``_nms_post_max_size`` 300, ``_nms_iou_threshold`` 0.1, ``_nms_score_threshold`` 0.05);
* ``infer_gpu(example, preds_dict)`` (:26-138; ``infer_torch`` is the same function with the decode on the device) --
  returns the list of annotation dicts (``name``, ``location``, ``dimensions``, ``rotation_y``, ``score`` + the empty
  KITTI fields of ``get_start_result_anno`` :724-737).

The reference walks the classes of ONE frame in Python with ~10 ``torch.cuda.synchronize()`` per class, copies the top-k
rows to the host, decodes in numpy and runs the numba NMS.  Here a batch of frames x classes goes through six launches
on the device (``PostProcessor``), and the only synchronisation is the final read-back of the detections; a batch of
more than one frame returns one annotation dict per frame (the reference handles batch 1 only: ``squeeze(0)``).

Order of equal scores: see include/pillarfront.h (pf_postprocess).  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


def _torch():
    import torch
    return torch


def get_start_result_anno():
    """framework/inference.py:724-737."""
    return {
        'name': np.array([]), 'truncated': np.array([]), 'occluded': np.array([]), 'alpha': np.array([]),
        'bbox': np.zeros([0, 4]), 'dimensions': np.zeros([0, 3]), 'location': np.zeros([0, 3]),
        'rotation_y': np.array([]), 'score': np.array([]),
    }


class PostProcessor:
    """Device-resident post-processing of ``max_frames`` frames: owns the anchors, the scratch and the output buffers.

    ``__call__(cls_preds, box_preds, dir_preds, anchors_mask)`` with CUDA fp32 tensors (B, A, n_ch) / (B, A, 7) /
    (B, A, 2) and a one-byte mask (B, A) or None -> ``(det_boxes (B, n_classes * post_max, 7), det_scores, det_class,
    det_anchor, det_num (B,))``, views of this object's buffers, valid until the next call; nothing synchronises."""

    def __init__(self, anchors, class_ranges, center_limit, max_frames=1, n_cls_channels=1, pre_max_size=1000,
                 post_max_size=300, iou_threshold=0.1, score_threshold=0.05, device=0):
        torch = _torch()
        if not torch.cuda.is_available():
            raise _lib.PillarFrontError("no CUDA device: the post-processing has no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device("cuda", device) if not isinstance(device, torch.device) else device
        a = np.ascontiguousarray(np.asarray(anchors, dtype=np.float32).reshape(-1, 7))
        self.n_anchors = int(a.shape[0])
        self.anchors = torch.from_numpy(a).to(self.device)
        ranges = [[int(lo), int(hi)] for lo, hi in class_ranges]
        self.class_ranges = ranges
        self.max_frames = int(max_frames)
        self.desc = _lib.make_post_desc(ranges, n_cls_channels, pre_max_size, post_max_size, score_threshold, iou_threshold,
                                        center_limit)
        need = int(self.lib.pf_post_scratch_bytes(C.byref(self.desc), self.max_frames, self.n_anchors))
        if need == 0:
            raise _lib.PillarFrontError("pf_post_scratch_bytes: " + self.lib.pf_last_error_string().decode())
        self.scratch = torch.empty((need,), dtype=torch.uint8, device=self.device)
        rows = len(ranges) * int(post_max_size)
        B = self.max_frames
        self.det_boxes = torch.empty((B, rows, 7), dtype=torch.float32, device=self.device)
        self.det_scores = torch.empty((B, rows), dtype=torch.float32, device=self.device)
        self.det_class = torch.empty((B, rows), dtype=torch.int32, device=self.device)
        self.det_anchor = torch.empty((B, rows), dtype=torch.int32, device=self.device)
        self.det_num = torch.zeros((B,), dtype=torch.int32, device=self.device)

    def _check(self, t, shape, what):
        torch = _torch()
        if not (hasattr(t, "is_cuda") and t.is_cuda) or t.dtype != torch.float32:
            raise TypeError(f"{what} must be a CUDA float32 tensor")
        if tuple(t.shape) != shape:
            raise ValueError(f"{what} must have shape {shape}, got {tuple(t.shape)}")
        return t.contiguous()

    def __call__(self, cls_preds, box_preds, dir_preds, anchors_mask=None):
        torch = _torch()
        B = int(cls_preds.shape[0])
        if B < 1 or B > self.max_frames:
            raise ValueError(f"batch of {B} frames outside [1, {self.max_frames}]")
        A, nch = self.n_anchors, int(self.desc.n_cls_channels)
        cls_preds = self._check(cls_preds, (B, A, nch), "cls_preds")
        box_preds = self._check(box_preds, (B, A, 7), "box_preds")
        dir_preds = self._check(dir_preds, (B, A, 2), "dir_preds")
        mask_ptr = None
        if anchors_mask is not None:
            if not anchors_mask.is_cuda or anchors_mask.element_size() != 1 or tuple(anchors_mask.shape) != (B, A):
                raise ValueError(f"anchors_mask must be a CUDA bool / uint8 tensor of shape {(B, A)}")
            anchors_mask = anchors_mask.contiguous()
            mask_ptr = anchors_mask.data_ptr()
        with _lib.on_device(self.device):
            _lib.check(self.lib.pf_postprocess(
                C.byref(self.desc), cls_preds.data_ptr(), box_preds.data_ptr(), dir_preds.data_ptr(), mask_ptr,
                self.anchors.data_ptr(), B, A, self.scratch.data_ptr(), self.det_boxes.data_ptr(),
                self.det_scores.data_ptr(), self.det_class.data_ptr(), self.det_anchor.data_ptr(), self.det_num.data_ptr(),
                _lib.raw_stream(self.device.index)), "pf_postprocess")
        return self.det_boxes[:B], self.det_scores[:B], self.det_class[:B], self.det_anchor[:B], self.det_num[:B]

    def region(self, name, n_frames):
        """Typed view of a scratch region of the last call over ``n_frames`` frames (tests / diagnostics)."""
        torch = _torch()
        idx, dtype, width = _lib.POST_REGIONS[name]
        offs = (C.c_size_t * len(_lib.POST_REGION_NAMES))()
        total = int(self.lib.pf_post_scratch_layout(C.byref(self.desc), int(n_frames), self.n_anchors, offs))
        end = int(offs[idx + 1]) if idx + 1 < len(_lib.POST_REGION_NAMES) else total
        raw = self.scratch[int(offs[idx]):end]
        t = raw.view(getattr(torch, dtype))
        if width > 1:       # regions are padded to 256 bytes: drop the partial row at the end
            t = t[:t.numel() // width * width].view(-1, width)
        return t


class Inference:
    """Drop-in for framework/inference.py:Inference (constructor :10-24, infer_gpu :26-138)."""

    def __init__(self, config, anchor_assigner, max_frames=None):
        torch = _torch()
        dev = config['device']                      # an index (configs/*.json), "cuda:0", or a torch.device (ros_node.py:37-38)
        name = str(dev)
        if isinstance(dev, int) or name.isdigit():
            index = int(name)
        elif name.startswith("cuda"):
            index = int(name.split(":")[1]) if ":" in name else 0
        else:
            raise _lib.PillarFrontError(f"config['device'] = {dev!r}: the post-processing has no CPU fallback")
        self.device = torch.device("cuda", index)
        self._anchors_np = np.ascontiguousarray(anchor_assigner.anchors, np.float32).reshape(-1, 7)
        self._nms_pre_max_size = 1000
        self._nms_post_max_size = 300
        self._nms_iou_threshold = 0.1
        self._box_code_size = 7
        self._num_class = 1
        self._use_direction_classifier = True
        self._nms_score_threshold = 0.05
        self.center_limit = config['center_limit']
        self.detect_class = np.array(config['detect_class'])
        self.class_masks = anchor_assigner.class_masks
        self._max_frames = int(max_frames if max_frames is not None else config.get('batch_size', 1) or 1)
        self._post = None
        self._post_key = None

    @property
    def anchors(self):
        return self._processor().anchors

    def _processor(self, n_frames=1, n_ch=1):
        key = (self._nms_pre_max_size, self._nms_post_max_size, float(self._nms_iou_threshold),
               float(self._nms_score_threshold), tuple(float(x) for x in self.center_limit), n_ch,
               max(self._max_frames, n_frames))
        if self._post is None or self._post_key != key:
            ranges = [self.class_masks[c] for c in self.class_masks]
            self._post = PostProcessor(self._anchors_np, ranges, self.center_limit, max_frames=key[-1], n_cls_channels=n_ch,
                                       pre_max_size=self._nms_pre_max_size, post_max_size=self._nms_post_max_size,
                                       iou_threshold=self._nms_iou_threshold, score_threshold=self._nms_score_threshold,
                                       device=self.device)
            self._post_key = key
        return self._post

    def infer_device(self, example, preds_dict):
        """The device half: returns the PostProcessor outputs for the batch (no synchronisation)."""
        torch = _torch()
        cls = preds_dict["cls_preds"]
        box = preds_dict["box_preds"]
        dirp = preds_dict["dir_preds"]
        if cls.dim() == 2:
            cls, box, dirp = cls.unsqueeze(0), box.unsqueeze(0), dirp.unsqueeze(0)
        B = int(cls.shape[0])
        mask = example["anchors_mask"] if example is not None and "anchors_mask" in example else None
        if mask is not None:
            if isinstance(mask, np.ndarray):
                mask = torch.from_numpy(np.ascontiguousarray(mask))
            mask = mask.to(self.device).reshape(B, -1)
            if mask.dtype not in (torch.bool, torch.uint8):
                mask = mask != 0
        post = self._processor(B, int(cls.shape[-1]))
        return post(cls.to(self.device, torch.float32), box.to(self.device, torch.float32),
                    dirp.to(self.device, torch.float32), mask)

    def infer_gpu(self, example, preds_dict):
        boxes, scores, klass, _, num = self.infer_device(example, preds_dict)
        num = num.cpu().numpy()                                   # the one synchronisation
        m = int(num.max()) if num.size else 0
        boxes = boxes[:, :m].cpu().numpy()
        scores = scores[:, :m].cpu().numpy()
        klass = klass[:, :m].cpu().numpy()
        names = list(self.class_masks.keys())
        annos = []
        for f in range(len(num)):
            n = int(num[f])
            anno = get_start_result_anno()
            if n > 0:
                b = boxes[f, :n]
                anno["name"] = np.array([names[k] for k in klass[f, :n]], dtype='<U10')
                anno["location"] = b[:, :3].copy()
                anno["dimensions"] = b[:, 3:6].copy()
                anno["rotation_y"] = b[:, 6].copy()
                anno["score"] = scores[f, :n].copy()
            annos.append(anno)
        return annos

    infer_torch = infer_gpu     # :140-256, the same loop with the decode in torch
    infer = infer_gpu           # what ros_node.py:75 calls (the reference's class has no such method any more)
