"""Single-frame inference path as ONE CUDA graph: pinned staging -> H2D -> voxelise + decorate + fused PFN -> canvas
fill -> D2H of the voxel count.  Used by ``VoxelGenerator.generate_and_scatter`` (the ``ros_node.py`` / ``train.py
infer`` loop of the reference: one frame at a time, framework/dataset.py:213-229, ros_node.py:57-75)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .frontend import PillarFrontEnd


class FusedFramePath:
    def __init__(self, gen, cap: int, folded, n_channels: int = 64):
        cfg = dict(detection_range=[float(x) for x in gen.detection_range], voxel_size=gen._voxel_size_cfg,
                   max_voxels=gen.max_voxels, max_num_points=gen.max_num_points, num_point_features=4)
        self.cap = cap
        self.fe = PillarFrontEnd(cfg, max_batch=1, max_points_per_frame=cap, device=gen._device_index,
                                 n_channels=n_channels, want_voxels=False, geometry=gen.geometry)
        self.dev = self.fe.device
        self.folded = folded
        with torch.cuda.device(self.dev):
            self.h_pts = torch.empty((cap, 4), dtype=torch.float32).pin_memory()
            self.h_offs = torch.zeros((2,), dtype=torch.int32).pin_memory()
            self.h_vnum = torch.zeros((1,), dtype=torch.int32).pin_memory()
            self.np_pts, self.np_offs, self.np_vnum = self.h_pts.numpy(), self.h_offs.numpy(), self.h_vnum.numpy()
            self.d_pts = torch.empty((cap, 4), dtype=torch.float32, device=self.dev)
            self.d_offs = torch.zeros((2,), dtype=torch.int32, device=self.dev)
            self.stream = torch.cuda.Stream(self.dev)
            self.event = torch.cuda.Event()
            self.event.record(self.stream)            # a torch event owns its cudaEvent_t from the first record on
            self.stream.synchronize()
        self.ev_done = self.event.cuda_event
        self.p_stage, self.p_offs, self.p_vnum = self.h_pts.data_ptr(), self.h_offs.data_ptr(), self.h_vnum.data_ptr()
        self.c_v = C.c_int32(0)
        self.graphs = {}                              # staged rows -> (CUDAGraph, cudaGraphExec_t)

    def _body(self, rows: int):
        if rows:                                      # rows == 0: pf_frame_run uploads the points in front of the graph
            self.d_pts[:rows].copy_(self.h_pts[:rows], non_blocking=True)
        self.d_offs.copy_(self.h_offs, non_blocking=True)
        feats = self.fe.voxelize_pfn(self.d_pts, self.d_offs, self.folded, n_frames=1, max_frame_points=self.cap)
        self.fe.scatter(feats, n_frames=1)
        self.h_vnum.copy_(self.fe.voxel_num, non_blocking=True)

    def _graph(self, rows: int):
        g = self.graphs.get(rows)
        if g is None:
            torch.cuda.synchronize(self.dev)
            with torch.cuda.device(self.dev), torch.cuda.stream(self.stream), torch.no_grad():
                self._body(rows)
                self.stream.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.stream):
                    self._body(rows)
            torch.cuda.synchronize(self.dev)
            g = (g, g.raw_cuda_graph_exec())
            self.graphs[rows] = g
        return g

    def run(self, points, n: int, threads: int):
        from .voxel_generator import _CHUNKED_UPLOAD, _raw_stream
        src = points if points.flags.c_contiguous else points.copy()
        rows = 0 if _CHUNKED_UPLOAD else min(self.cap, max(4096, (n + 4095) // 4096 * 4096))
        _, exec_ptr = self._graph(rows)
        # stage, replay on the caller's current stream (ordered behind the consumer of the previous canvas, in front of the
        # consumer of this one), wait, read V: one foreign call
        idx = self.dev.index
        rc = self.fe.handle.lib.pf_frame_run(idx, exec_ptr, _raw_stream(idx), self.ev_done, None, self.p_stage,
                                             src.ctypes.data if n else None, src.nbytes if n else 0, threads,
                                             self.d_pts.data_ptr() if _CHUNKED_UPLOAD else None, self.p_offs, n,
                                             self.p_vnum, C.byref(self.c_v))
        if rc:
            _lib.check(rc, "pf_frame_run")
        v = self.c_v.value
        return self.fe.canvas, self.fe.coors[:v], v

    def close(self):
        self.graphs.clear()
        self.fe.close()
