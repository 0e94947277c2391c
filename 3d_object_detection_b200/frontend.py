"""Batched, device-resident front end: voxelise -> decorate -> (PFN) -> BEV scatter for B frames per step.

One :class:`PillarFrontEnd` owns one workspace handle, one set of preallocated output tensors and (optionally)
one CUDA graph of the whole step, for one GPU and one stream.  Frames are independent, so multi-GPU use is
"one PillarFrontEnd per rank, frames sharded across ranks, no collective" (see sharding.py).

Per step (all on the current stream, no host synchronisation):

    pf_voxelize_decorate(points, frame_offsets) -> voxels, coors, num, voxel_num, features   (6-7 kernels)
    pillar features (B*max_voxels, C) = pfn(features)  or a caller-supplied tensor            (not on the path)
    pf_scatter_from_map(pillar features)              -> canvas (B, C, nx, ny)                (1 kernel)

The voxeliser's own cell grid doubles as the cell -> pillar map of the scatter, so coordinates are never
re-indexed and the canvas is written exactly once.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .config import derive_grid
from .pillar_ops import decorate_constants


class PillarFrontEnd:
    N_KERNELS_SCATTER = 1    # canvas_fill

    @staticmethod
    def n_kernels_voxelize(max_frame_points: int) -> int:
        """Launches of one pf_voxelize_decorate call: quantize (two index ranges once a frame has >= 32768 points,
        csrc/pf_voxelize.cu), assign, tile_rank, column_scan, slot, gather(+decorate)."""
        return (2 if max_frame_points >= 32 * 1024 else 1) + 5

    def __init__(self, config, max_batch, max_points_per_frame, device=0, n_channels=64, want_voxels=True,
                 compact=False, geometry=None):
        if not torch.cuda.is_available():
            raise _lib.PillarFrontError("no CUDA device: the pillar front end has no CPU fallback")
        cfg = dict(config)
        # `geometry`: an already derived grid (derive_grid snaps the range, so deriving twice from a snapped range is
        # not guaranteed to be the identity in fp32)
        self.geometry = geometry if geometry is not None else derive_grid(cfg)
        g = self.geometry
        self.device = torch.device("cuda", int(device))
        self.B = int(max_batch)
        self.max_points_per_frame = int(max_points_per_frame)
        self.max_voxels = int(cfg["max_voxels"])
        self.P = int(cfg["max_num_points"])
        self.F = int(cfg.get("num_point_features", 4))
        self.C = int(n_channels)
        self.nx, self.ny, self.nz = (int(x) for x in g.grid_size)
        if self.nz != 1:
            raise _lib.PillarFrontError("PillarFrontEnd needs nz == 1 (pillars); use VoxelGenerator + PointPillarsScatter")
        self.compact = bool(compact)
        self.consts = decorate_constants(cfg["voxel_size"], g.offset)
        self.handle = _lib.Handle(g, self.max_voxels, self.P, self.F, self.B, self.max_points_per_frame,
                                  self.device.index)
        rows = self.B * self.max_voxels
        dev = self.device
        self.voxels = torch.empty((rows, self.P, self.F), dtype=torch.float32, device=dev) if want_voxels else None
        self.coors = torch.empty((rows, 4), dtype=torch.int32, device=dev)
        self.num_points = torch.empty((rows,), dtype=torch.int32, device=dev)
        self.voxel_num = torch.zeros((self.B,), dtype=torch.int32, device=dev)
        self.features = torch.empty((rows, self.P, self.F + 5), dtype=torch.float32, device=dev)
        self.canvas = torch.empty((self.B, self.C, self.nx, self.ny), dtype=torch.float32, device=dev)
        self.pillar_feats = None   # (rows, 64), allocated by the first voxelize_pfn
        self._graph = None
        self._g_inputs = None

    # ---- the two stages --------------------------------------------------------------------------
    def voxelize_decorate(self, points, frame_offsets, n_frames=None, max_frame_points=None, keep_cell_map=True):
        """points: CUDA fp32 (sum N_f, F); frame_offsets: CUDA int32 (n_frames+1,).  With ``keep_cell_map`` (default)
        the call must be followed by exactly one ``scatter`` (or ``discard_cell_map``) before the next voxelise."""
        n_frames = self.B if n_frames is None else int(n_frames)
        mfp = self.max_points_per_frame if max_frame_points is None else int(max_frame_points)
        h = self.handle
        vx, vy, xo, yo = self.consts
        stream = _lib.raw_stream(self.device.index)
        _lib.check(h.lib.pf_voxelize_decorate(
            h.ptr, points.data_ptr(), frame_offsets.data_ptr(), n_frames, mfp, int(points.shape[1]),
            1 if self.compact else 0, 4, vx, vy, xo, yo, 1 if keep_cell_map else 0,
            self.voxels.data_ptr() if self.voxels is not None else None, self.coors.data_ptr(),
            self.num_points.data_ptr(), self.voxel_num.data_ptr(), self.features.data_ptr(), stream),
            "pf_voxelize_decorate")

    def discard_cell_map(self):
        """Drop the cell map a voxelise call kept when no scatter follows it."""
        h = self.handle
        _lib.check(h.lib.pf_discard_cell_map(h.ptr, _lib.raw_stream(self.device.index)), "pf_discard_cell_map")

    def voxelize_pfn(self, points, frame_offsets, folded, n_frames=None, max_frame_points=None, want_features=False):
        """Fused voxelise + decorate + PFN (pf_voxelize_pfn): fills voxels/coors/num/voxel_num and ``self.pillar_feats``
        (rows, 64); ``folded`` = (weight (64,9), scale, shift) from pillar_ops.fold_pfn / PointNet.folded()."""
        n_frames = self.B if n_frames is None else int(n_frames)
        mfp = self.max_points_per_frame if max_frame_points is None else int(max_frame_points)
        if self.pillar_feats is None:
            self.pillar_feats = torch.empty((self.B * self.max_voxels, 64), dtype=torch.float32, device=self.device)
        w, sc, sh = folded
        h = self.handle
        vx, vy, xo, yo = self.consts
        stream = _lib.raw_stream(self.device.index)
        _lib.check(h.lib.pf_voxelize_pfn(
            h.ptr, points.data_ptr(), frame_offsets.data_ptr(), n_frames, mfp, int(points.shape[1]),
            1 if self.compact else 0, 4, vx, vy, xo, yo, 1, w.data_ptr(), sc.data_ptr(), sh.data_ptr(), 64,
            self.voxels.data_ptr() if self.voxels is not None else None, self.coors.data_ptr(),
            self.num_points.data_ptr(), self.voxel_num.data_ptr(),
            self.features.data_ptr() if want_features else None, self.pillar_feats.data_ptr(), stream),
            "pf_voxelize_pfn")
        return self.pillar_feats

    def scatter(self, pillar_feats, n_frames=None):
        """pillar_feats: CUDA fp32 (rows, C) row-major, rows laid out like the voxel outputs."""
        n_frames = self.B if n_frames is None else int(n_frames)
        h = self.handle
        stream = _lib.raw_stream(self.device.index)
        _lib.check(h.lib.pf_scatter_from_map(h.ptr, pillar_feats.data_ptr(), int(pillar_feats.stride(0)),
                                             int(pillar_feats.stride(1)), n_frames, 1 if self.compact else 0,
                                             self.C, self.canvas.data_ptr(), stream), "pf_scatter_from_map")
        return self.canvas[:n_frames]

    def step(self, points, frame_offsets, pillar_feats=None, pfn=None, n_frames=None, max_frame_points=None):
        """One pass of the hot path over one batch.  Exactly one of pillar_feats / pfn feeds the scatter."""
        fused = (pfn is not None and hasattr(pfn, "folded") and self.F == 4 and
                 (pfn.fusable() if hasattr(pfn, "fusable") else not getattr(pfn, "training", True)))
        if fused:   # PointNet in eval mode: decorate + conv + BN + ReLU + max inside the gather kernel
            return self.scatter(self.voxelize_pfn(points, frame_offsets, pfn.folded(), n_frames, max_frame_points),
                                n_frames)
        self.voxelize_decorate(points, frame_offsets, n_frames, max_frame_points)
        if pfn is not None:   # PointNet in training mode -> its torch tail; any other callable maps features -> (rows, C)
            pillar_feats = pfn.features_to_pillars(self.features) if hasattr(pfn, "features_to_pillars") else pfn(self.features)
        if pillar_feats is None:
            raise ValueError("step needs pillar_feats or pfn")
        return self.scatter(pillar_feats, n_frames)

    def anchor_mask(self, anchors, n_frames=None):
        """Anchor occupancy mask (n_frames, A) torch.bool of the frames last voxelised, straight from the device
        coordinates (row N2; ``anchors`` is an anchor_mask.AnchorMask built for this grid)."""
        n_frames = self.B if n_frames is None else int(n_frames)
        rows = n_frames * self.max_voxels
        return anchors.create_mask_batch(self.coors[:rows], self.voxel_num, n_frames,
                                         0 if self.compact else self.max_voxels)

    # ---- CUDA graph of a fixed-shape step --------------------------------------------------------
    def capture(self, points, frame_offsets, pillar_feats=None, pfn=None, chain_priority=False):
        """Capture step() for these (static) tensors; replay() re-runs it on whatever they then contain.  With ``pfn``
        (a PointNet in eval mode) the fused voxelise + decorate + PFN kernel is captured; its folded weights are
        read at capture time.

        ``chain_priority``: capture the voxelise chain on a HIGH-priority stream and the canvas fill on a normal one
        (fork / join inside the graph).  When several front ends replay on different streams -- batch k+1's chain while
        batch k's fill is still streaming out -- the latency-bound chain kernels are then scheduled ahead of the fill's
        thousands of CTAs instead of queueing behind them."""
        self._g_inputs = (points, frame_offsets, pillar_feats, pfn)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):
                self.step(points, frame_offsets, pillar_feats=pillar_feats, pfn=pfn)
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        if not chain_priority:
            with torch.cuda.graph(g), torch.no_grad():
                self.step(points, frame_offsets, pillar_feats=pillar_feats, pfn=pfn)
        else:
            lo, hi_prio = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
            hi = torch.cuda.Stream(self.device, priority=hi_prio)
            cap = torch.cuda.Stream(self.device)
            fused = (pfn is not None and hasattr(pfn, "folded") and self.F == 4 and
                     (pfn.fusable() if hasattr(pfn, "fusable") else not getattr(pfn, "training", True)))
            with torch.cuda.graph(g, stream=cap), torch.no_grad():
                hi.wait_stream(cap)                                   # fork
                with torch.cuda.stream(hi):
                    if fused:
                        feats = self.voxelize_pfn(points, frame_offsets, pfn.folded())
                    else:
                        self.voxelize_decorate(points, frame_offsets)
                        feats = pillar_feats
                        if pfn is not None:
                            feats = pfn.features_to_pillars(self.features) if hasattr(pfn, "features_to_pillars") else pfn(self.features)
                cap.wait_stream(hi)                                   # join
                self.scatter(feats)
        self._graph = g
        return g

    def replay(self):
        self._graph.replay()
        return self.canvas

    # ---- host-side views -------------------------------------------------------------------------
    def frame_outputs(self, f, counts=None):
        """Slices of frame f in the slab layout (synchronises to read voxel_num unless counts is given)."""
        if self.compact:
            raise ValueError("frame_outputs is for the slab layout")
        counts = self.voxel_num.cpu().numpy() if counts is None else counts
        v = int(counts[f])
        lo = f * self.max_voxels
        sl = slice(lo, lo + v)
        return (self.voxels[sl] if self.voxels is not None else None, self.coors[sl], self.num_points[sl],
                self.features[sl])

    def bytes_algorithmic(self, n_points_total, voxel_counts):
        """API-visible bytes of one step: each tensor once, scratch excluded."""
        v = int(np.sum(voxel_counts))
        b = len(voxel_counts)
        P, F, C = self.P, self.F, self.C
        vox = 4 * F * n_points_total + 4 * F * P * v + 12 * v + 4 * v
        dec = 4 * (F + 5) * P * v
        scat = 4 * C * v + 12 * v + 4 * C * self.nx * self.ny * b
        return {"voxelize": vox, "decorate_out": dec, "scatter": scat, "e2e": vox + dec + scat}

    def close(self):
        self.handle.close()
