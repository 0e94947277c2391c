"""Anchor occupancy mask ("next" row N2) backed by ``pf_anchor_mask`` (csrc/pf_anchor.cu).

Mirrors, name for name, what the reference's data path calls once per frame:

* ``get_anchor_coor(anchors_bv, stride, offset, grid_size)`` -- framework/box_np_ops.py:288-304, host-side and
  init-time (numpy fp32 arithmetic, the same operations in the same order as the numba loop);
* ``fused_get_anchors_mask_gpu(coors, shape, anchor_coors_cuda, anchors_mask_cuda)`` -- box_np_ops.py:236-257;
* ``AnchorMask.create_mask(coors, grid_size, voxel_size, offset, gpu=True)`` -- the method of
  ``AnchorAssigner`` (framework/anchor_assigner.py:322-335) with the two device buffers its ``__init__`` makes
  (:295-298).

The reference zero-fills and uploads an (nx, ny) fp32 map per frame and scans it with one thread per row /
column; here the pillar-count table is int32 (exact: counts <= max_voxels < 2^24), stays on the device, and a
whole batch is handled by four launches.  ``create_mask_batch`` / ``PillarFrontEnd.anchor_mask`` take the device
coordinates straight from the voxeliser, so nothing crosses PCIe.

No CPU fallback: ``gpu=False`` raises (the reference's numpy branch is restated in ``oracle/`` for the tests only).
"""
from __future__ import annotations

import numpy as np

from . import _lib


def _torch():
    import torch
    return torch


def get_anchor_coor(anchors_bv, stride, offset, grid_size):
    """(A, 4) int32 ``(minx, miny, maxx, maxy)`` cell rectangle of every bird's-eye anchor box; clipped to the grid
    on the low side of the minima and the high side of the maxima only, exactly like box_np_ops.py:296-303."""
    bv = np.asarray(anchors_bv)
    stride = np.asarray(stride)
    offset = np.asarray(offset)
    ret = np.empty(bv.shape, dtype=np.int32)
    for k in range(4):
        q = np.floor((bv[:, k] - offset[k & 1]) / stride[k & 1])      # dtype follows the inputs, as in numba
        ret[:, k] = q.astype(np.int32)
    np.maximum(ret[:, 0], 0, out=ret[:, 0])
    np.maximum(ret[:, 1], 0, out=ret[:, 1])
    np.minimum(ret[:, 2], int(grid_size[0]) - 1, out=ret[:, 2])
    np.minimum(ret[:, 3], int(grid_size[1]) - 1, out=ret[:, 3])
    return ret


def _device_coors(coors, device):
    torch = _torch()
    if isinstance(coors, np.ndarray):
        if coors.dtype != np.int32:
            coors = coors.astype(np.int32)
        coors = torch.from_numpy(np.ascontiguousarray(coors)).to(device, non_blocking=True)
    else:
        coors = coors.to(device=device, dtype=torch.int32).contiguous()
    if coors.dim() != 2 or coors.shape[1] < 2:
        raise ValueError(f"coors must be (V, >=2), got {tuple(coors.shape)}")
    return coors


def anchor_mask_device(coors, shape, anchor_coors_cuda, anchors_mask_cuda=None, voxel_num=None, n_frames=1,
                       rows_per_frame=0, integral=None):
    """Device-to-device form: returns (mask (n_frames, A) torch.bool CUDA, integral scratch (n_frames, nx, ny) int32
    CUDA; ``integral_table`` gives the typed view of it).  ``coors`` CUDA int32 (rows, >=2); layouts as documented at
    pf_anchor_mask in include/pillarfront.h."""
    torch = _torch()
    if not torch.cuda.is_available():
        raise _lib.PillarFrontError("no CUDA device: the anchor mask has no CPU fallback")
    lib = _lib.load()
    nx, ny = int(shape[0]), int(shape[1])
    dev = anchor_coors_cuda.device
    A = int(anchor_coors_cuda.shape[0])
    if anchor_coors_cuda.dtype != torch.int32 or not anchor_coors_cuda.is_contiguous() or anchor_coors_cuda.shape[1] != 4:
        raise ValueError("anchor_coors_cuda must be a contiguous CUDA int32 (A, 4) tensor")
    if integral is None:
        integral = torch.empty((n_frames, nx, ny), dtype=torch.int32, device=dev)
    if anchors_mask_cuda is None:
        anchors_mask_cuda = torch.empty((n_frames, A), dtype=torch.bool, device=dev)
    if anchors_mask_cuda.numel() < n_frames * A or anchors_mask_cuda.element_size() != 1:
        raise ValueError("anchors_mask_cuda must hold n_frames * A one-byte elements")
    if coors.dtype != torch.int32 or coors.dim() != 2 or coors.shape[1] < 2 or (coors.numel() and coors.stride(1) != 1):
        raise ValueError("coors must be a CUDA int32 (rows, >=2) tensor with unit column stride")
    row_stride = int(coors.stride(0)) if coors.shape[0] > 1 else int(coors.shape[1])
    stream = _lib.raw_stream(dev.index)
    with _lib.on_device(dev):
        _lib.check(lib.pf_anchor_mask(coors.data_ptr() if coors.numel() else None, row_stride,
                                      voxel_num.data_ptr() if voxel_num is not None else None, int(coors.shape[0]),
                                      int(rows_per_frame), int(n_frames), nx, ny, anchor_coors_cuda.data_ptr(), A,
                                      integral.data_ptr(), anchors_mask_cuda.data_ptr(), stream), "pf_anchor_mask")
    return anchors_mask_cuda, integral


def integral_table(integral, n_rows, rows_per_frame=0):
    """numpy view of the running-sum table pf_anchor_mask left in its scratch: uint16 when a frame cannot hold more
    than 65535 pillars, else int32 (the rule of csrc/pf_anchor.cu launch_anchor_mask)."""
    n_frames, nx, ny = integral.shape
    raw = integral.cpu().numpy().reshape(-1)
    bound = rows_per_frame if rows_per_frame > 0 else n_rows
    if bound <= 65535:
        return raw.view(np.uint16)[: n_frames * nx * ny].reshape(n_frames, nx, ny)
    return raw.reshape(n_frames, nx, ny)


def fused_get_anchors_mask_gpu(coors, shape, anchor_coors_cuda, anchors_mask_cuda):
    """Same signature and return value (numpy bool (A,)) as box_np_ops.fused_get_anchors_mask_gpu; ``coors`` is the
    (V, 3) int32 array of VoxelGenerator.generate (numpy or torch), the two ``*_cuda`` arguments torch CUDA tensors."""
    coors_d = _device_coors(coors, anchor_coors_cuda.device)
    mask, _ = anchor_mask_device(coors_d, shape, anchor_coors_cuda, anchors_mask_cuda, None, 1, 0)
    return mask.reshape(-1)[: anchor_coors_cuda.shape[0]].to(_torch().bool).cpu().numpy()


class AnchorMask:
    """The mask-making part of AnchorAssigner: holds ``anchors_coors_cuda`` / ``anchors_mask_cuda`` like
    framework/anchor_assigner.py:295-298 and offers the reference's ``create_mask``."""

    def __init__(self, anchors_bv, config=None, voxel_size=None, offset=None, grid_size=None, device=0):
        torch = _torch()
        if not torch.cuda.is_available():
            raise _lib.PillarFrontError("no CUDA device: the anchor mask has no CPU fallback")
        if config is not None:
            voxel_size = np.array(config["voxel_size"], dtype=np.float32)
            offset = np.array(config["detection_offset"], dtype=np.float32)
            grid_size = config["grid_size"]
        self.grid_size = np.asarray(grid_size)
        self.anchors_bv = np.asarray(anchors_bv)
        self.anchors_coors = get_anchor_coor(self.anchors_bv, voxel_size, offset, self.grid_size)
        self.device = torch.device("cuda", int(device))
        self.anchors_coors_cuda = torch.from_numpy(self.anchors_coors).to(self.device)
        self.anchors_mask_cuda = torch.zeros((self.anchors_coors.shape[0],), dtype=torch.bool, device=self.device)
        self._scratch = {}

    @property
    def n_anchors(self):
        return int(self.anchors_coors.shape[0])

    def create_mask(self, coors, grid_size, voxel_size, offset, gpu=True):
        if not gpu:
            raise _lib.PillarFrontError("create_mask(gpu=False): this library has no CPU path")
        return fused_get_anchors_mask_gpu(coors, tuple(grid_size[:-1]), self.anchors_coors_cuda, self.anchors_mask_cuda)

    def create_mask_batch(self, coors, voxel_num=None, n_frames=1, rows_per_frame=0):
        """Device in, device out: (n_frames, A) torch.bool; buffers are reused between calls of the same batch size."""
        torch = _torch()
        key = int(n_frames)
        if key not in self._scratch:
            nx, ny = int(self.grid_size[0]), int(self.grid_size[1])
            self._scratch[key] = (torch.empty((key, self.n_anchors), dtype=torch.bool, device=self.device),
                                  torch.empty((key, nx, ny), dtype=torch.int32, device=self.device))
        mask, integ = self._scratch[key]
        anchor_mask_device(coors, self.grid_size[:2], self.anchors_coors_cuda, mask, voxel_num, n_frames,
                           rows_per_frame, integ)
        return mask
