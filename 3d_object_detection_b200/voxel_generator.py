"""Drop-in for the reference's ``framework/voxel_generator.py`` backed by libpillarfront.so (CUDA, sm_100a).

Same constructor argument (the config dict), same attributes, same config side effects and
the same return values as ``VoxelGenerator.generate`` (framework/voxel_generator.py:28-40): ``voxels``
``(V, P, F)`` fp32, ``coors`` ``(V, 3)`` int32 in **x, y, z** order (F3) and ``num_points_per_voxel`` ``(V,)``
int32, bit-identical to the numba loop including its ``max_voxels`` early ``break`` (F5).

Differences, all deliberate (SURVEY.md §8b "Errors", "Threading"):

* points must be float32 ``(N, F>=3)``; anything else raises ``TypeError`` / ``ValueError`` (the reference
  would JIT a float64 specialisation).
* CUDA cannot be used from forked DataLoader workers: use ``num_workers = 0`` (what ntusl_20cm.json,
  ntusl_10cm.json and seven_20cm.json already set) or a ``spawn`` context.
* ``generate(points, device_outputs=True)`` (or CUDA tensor input) returns torch CUDA tensors and skips the
  device->host copy; ``example_convert_to_torch`` (framework/utils.py:7-20) passes them through unchanged.
* ``generate_batch`` voxelises a list of frames in one launch sequence and returns the collated layout of
  ``merge_second_batch`` (framework/utils.py:23-48): concatenated rows, frame id in the LAST coors column.

There is no CPU fallback: without a CUDA device ``generate`` raises.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from .config import derive_grid


def _torch():
    import torch
    return torch


class VoxelGenerator:
    _returns_voxel_num = False

    def __init__(self, config, device=None, max_points_per_frame=None):
        geom = derive_grid(config)
        self.geometry = geom
        self.voxel_size = geom.voxel_size
        self.detection_range = geom.detection_range
        self.offset = geom.offset
        self.grid_size = geom.grid_size
        self.max_num_points = int(config["max_num_points"])
        self.max_voxels = int(config["max_voxels"])
        self.num_point_features = int(config.get("num_point_features", 4))
        # the reference writes these four keys back; AnchorAssigner, the datasets and the networks read them
        # (framework/voxel_generator.py:23-26)
        config["detection_range"] = geom.detection_range
        config["detection_offset"] = geom.offset
        config["detection_range_diff"] = geom.range_diff
        config["grid_size"] = geom.grid_size
        dev = device if device is not None else config.get("device", 0)
        self._device_index = self._parse_device(dev)
        self._cap_hint = int(max_points_per_frame) if max_points_per_frame else 0
        self._handles = {}      # (max_frames, F) -> _lib.Handle
        self.voxel_time = 0.0   # kept for parity with InferData's timers (framework/dataset.py:208)

    # ------------------------------------------------------------------------------------------------
    @staticmethod
    def _parse_device(dev) -> int:
        if isinstance(dev, int):
            return dev
        s = str(dev)
        if s.startswith("cuda"):
            return int(s.split(":")[1]) if ":" in s else 0
        if s.isdigit():
            return int(s)
        raise ValueError(f"VoxelGenerator needs a CUDA device, got {dev!r} (no CPU fallback)")

    def _handle(self, n_frames: int, n_feat: int, max_pts: int) -> "_lib.Handle":
        torch = _torch()
        if not torch.cuda.is_available():
            raise _lib.PillarFrontError("no CUDA device: the pillar front end has no CPU fallback")
        key = n_feat
        h = self._handles.get(key)
        if h is None or h.max_frames < n_frames or h.max_points_per_frame < max_pts:
            frames = max(n_frames, h.max_frames if h else 1)
            cap = max(max_pts, self._cap_hint, 1)
            if h is not None:
                cap = max(cap, h.max_points_per_frame)
                if max_pts > h.max_points_per_frame:
                    cap = max(cap, int(1.5 * h.max_points_per_frame))
                torch.cuda.synchronize(self._device_index)
                h.close()
            h = _lib.Handle(self.geometry, self.max_voxels, self.max_num_points, n_feat, frames, cap,
                            self._device_index)
            self._handles[key] = h
        return h

    def _to_device(self, points):
        """Validate one frame and return a contiguous fp32 CUDA tensor (N, F)."""
        torch = _torch()
        if not torch.cuda.is_available():
            raise _lib.PillarFrontError("no CUDA device: the pillar front end has no CPU fallback")
        dev = torch.device("cuda", self._device_index)
        if isinstance(points, np.ndarray):
            if points.dtype != np.float32:
                raise TypeError(f"points must be float32, got {points.dtype}")
            if points.ndim != 2 or points.shape[1] < 3:
                raise ValueError(f"points must have shape (N, F>=3), got {points.shape}")
            t = torch.from_numpy(np.ascontiguousarray(points))
            return t.to(dev, non_blocking=True), False
        if isinstance(points, torch.Tensor):
            if points.dtype != torch.float32:
                raise TypeError(f"points must be float32, got {points.dtype}")
            if points.dim() != 2 or points.shape[1] < 3:
                raise ValueError(f"points must have shape (N, F>=3), got {tuple(points.shape)}")
            on_dev = points.is_cuda
            return points.contiguous().to(dev, non_blocking=True), on_dev
        raise TypeError(f"points must be a numpy array or a torch tensor, got {type(points).__name__}")

    # ------------------------------------------------------------------------------------------------
    def _run(self, frames_dev, compact: bool, coors_stride: int):
        torch = _torch()
        dev = torch.device("cuda", self._device_index)
        n_frames = len(frames_dev)
        n_feat = int(frames_dev[0].shape[1])
        if any(int(f.shape[1]) != n_feat for f in frames_dev):
            raise ValueError("all frames of a batch must have the same number of point features")
        if not 3 <= n_feat <= 16:
            raise ValueError(f"point features must be 3..16, got {n_feat}")
        sizes = [int(f.shape[0]) for f in frames_dev]
        max_pts = max(sizes + [1])
        h = self._handle(n_frames, n_feat, max_pts)
        with torch.cuda.device(dev):
            pts = frames_dev[0] if n_frames == 1 else torch.cat(frames_dev, dim=0)
            if pts.numel() == 0:
                pts = torch.zeros((1, n_feat), dtype=torch.float32, device=dev)
            offs = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32), device=dev)
            rows = n_frames * self.max_voxels
            voxels = torch.empty((rows, self.max_num_points, n_feat), dtype=torch.float32, device=dev)
            coors = torch.empty((rows, coors_stride), dtype=torch.int32, device=dev)
            num = torch.empty((rows,), dtype=torch.int32, device=dev)
            vnum = torch.empty((n_frames,), dtype=torch.int32, device=dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(h.lib.pf_voxelize(h.ptr, pts.data_ptr(), offs.data_ptr(), n_frames, max_pts, n_feat,
                                         1 if compact else 0, coors_stride, voxels.data_ptr(), coors.data_ptr(),
                                         num.data_ptr(), vnum.data_ptr(), stream), "pf_voxelize")
        return voxels, coors, num, vnum

    def generate(self, points, device_outputs=None):
        """One frame -> (voxels, coors, num_points_per_voxel)[, voxel_num]; numpy in -> numpy out."""
        import time
        t0 = time.time()
        dev_pts, was_cuda = self._to_device(points)
        if device_outputs is None:
            device_outputs = was_cuda
        voxels, coors, num, vnum = self._run([dev_pts], compact=False, coors_stride=3)
        v = int(vnum.item())                       # the only synchronisation: V is data dependent
        voxels, coors, num = voxels[:v], coors[:v], num[:v]
        if not device_outputs:
            voxels, coors, num = voxels.cpu().numpy(), coors.cpu().numpy(), num.cpu().numpy()
        self.voxel_time += time.time() - t0
        if self._returns_voxel_num:
            return voxels, coors, num, v
        return voxels, coors, num

    def generate_batch(self, frames, device_outputs=True):
        """List of frames -> dict in the collated layout of merge_second_batch (framework/utils.py:23-48)."""
        if len(frames) == 0:
            raise ValueError("generate_batch needs at least one frame")
        devs = [self._to_device(f)[0] for f in frames]
        stride = 4 if len(frames) > 1 else 3       # merge_second_batch pads the frame id only when B > 1
        voxels, coors, num, vnum = self._run(devs, compact=True, coors_stride=stride)
        counts = vnum.cpu().numpy()
        total = int(counts.sum())
        out = {"voxels": voxels[:total], "coordinates": coors[:total], "num_points_per_voxel": num[:total],
               "voxel_num": counts}
        if not device_outputs:
            out = {k: (v.cpu().numpy() if hasattr(v, "cpu") else v) for k, v in out.items()}
        return out

    def close(self):
        for h in self._handles.values():
            h.close()
        self._handles.clear()


class VoxelGenerator_trt(VoxelGenerator):
    """framework/voxel_generator.py:43-79: same, but generate() also returns voxel_num."""
    _returns_voxel_num = True
