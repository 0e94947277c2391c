"""Drop-in for the reference's ``framework/voxel_generator.py`` backed by libpillarfront.so (CUDA, sm_100a).

Same constructor argument (the config dict), same attributes, same config side effects and
the same return values as ``VoxelGenerator.generate`` (framework/voxel_generator.py:28-40): ``voxels``
``(V, P, F)`` fp32, ``coors`` ``(V, 3)`` int32 in **x, y, z** order (F3) and ``num_points_per_voxel`` ``(V,)``
int32, bit-identical to the numba loop including its ``max_voxels`` early ``break`` (F5).

The single-frame path is the reference's real operating point (every ``configs/*.json`` has ``batch_size: 1``;
``InferData.get``, framework/dataset.py:213-229; ros_node.py:70-75), so ``generate`` is built for latency:

* per instance and per ``(F, point-capacity bucket)`` one :class:`_FramePath` -- workspace handle, device buffers, a
  PINNED input staging buffer, and ONE CUDA graph holding "H2D of the staged frame + frame length, the voxelise
  kernels, D2H of the voxel count"; nothing is allocated, no ``torch.tensor(...)`` / ``.item()`` per call;
* ONE foreign call per frame (``pf_frame_run``): the numpy frame is copied into the staging buffer by a few host threads
  with non-temporal stores (a plain memcpy leaves dirty cache lines the copy engine then has to pull out of the CPU caches:
  +45 us on the H2D), the graph is replayed on the caller's current stream, the host waits on one event and reads ``V``
  from pinned memory; a second call (``pf_frame_copy_out``) copies exactly ``V`` rows out;
* numpy outputs land directly in a pinned block taken from a small recycling pool and are returned as views of it:
  a block is only handed out again once every array that was returned from it has been released (the reference's
  ownership contract -- callers may keep the arrays -- holds; nothing aliases a live result; an asynchronous copy a
  caller starts FROM such an array must have completed before it drops the array, which is how
  ``example_convert_to_torch``'s ``torch.as_tensor(..., device=)`` behaves);
* ``device_outputs=True`` (or CUDA tensor input) returns fresh torch CUDA tensors of exactly ``V`` rows;
* :meth:`VoxelGenerator.generate_and_scatter` goes one further for ``ros_node.py``-style loops: voxelise + decorate +
  fused PFN + canvas fill as one graph, one synchronisation per frame.

Differences, all deliberate (SURVEY.md §8b "Errors", "Threading"):

* points must be float32 ``(N, F>=3)``; anything else raises ``TypeError`` / ``ValueError`` (the reference
  would JIT a float64 specialisation).
* CUDA cannot be used from forked DataLoader workers: use ``num_workers = 0`` (what ntusl_20cm.json,
  ntusl_10cm.json and seven_20cm.json already set) or a ``spawn`` context.
* ``generate_batch`` voxelises a list of frames in one launch sequence and returns the collated layout of
  ``merge_second_batch`` (framework/utils.py:23-48): concatenated rows, frame id in the LAST coors column.

There is no CPU fallback: without a CUDA device ``generate`` raises.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
import time

import numpy as np

from . import _lib
from .config import derive_grid

_BUCKET = 32768          # point-capacity granularity of the single-frame graphs
# PF_FRAME_UPLOAD=chunked: pf_frame_run uploads the points itself, piece by piece as they are staged, in front of a graph
# without a points copy.  Measured and NOT the default (profiles/r2_frame_probe_upload.txt): the copy inside the graph wins
# by 3-10 us (20 cm 138 vs 142 us, nuScenes 243 vs 252, 10 cm 218 vs 222): four extra cudaMemcpyAsync calls and a graph
# launch behind them cost more than the staging time they hide.
_CHUNKED_UPLOAD = os.environ.get("PF_FRAME_UPLOAD", "graph") == "chunked"
_POOL_MAX = 8            # pinned output blocks kept per instance before falling back to pageable copies


def _torch():
    import torch
    return torch


_raw_stream = _lib.raw_stream


def _stage_threads() -> int:
    env = os.environ.get("PF_STAGE_THREADS", "")
    if env.isdigit() and int(env) >= 1:
        return min(16, int(env))
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return max(1, min(8, n // 2))     # streaming stores: 1.92 MB in 76 / 39 / 22 / 14 us with 1 / 2 / 4 / 8 threads


class _OutBlock:
    """One pinned host block for the numpy outputs of a frame; recycled once all views handed out from it died."""

    def __init__(self, torch, mv, P, F):
        self.tv = torch.empty((mv, P, F), dtype=torch.float32).pin_memory()
        self.tc = torch.empty((mv, 3), dtype=torch.int32).pin_memory()
        self.tn = torch.empty((mv,), dtype=torch.int32).pin_memory()
        self.nv, self.nc, self.nn = self.tv.numpy(), self.tc.numpy(), self.tn.numpy()
        self.p_v, self.p_c, self.p_n = self.tv.data_ptr(), self.tc.data_ptr(), self.tn.data_ptr()
        self._base = tuple(sys.getrefcount(a) for a in (self.nv, self.nc, self.nn))

    def free(self) -> bool:
        return tuple(sys.getrefcount(a) for a in (self.nv, self.nc, self.nn)) == self._base


class _FramePath:
    """Everything one single-frame call needs for a fixed (F, capacity): buffers, staging, graph."""

    def __init__(self, gen: "VoxelGenerator", n_feat: int, cap: int):
        torch = _torch()
        self.cap, self.F = cap, n_feat
        dev = torch.device("cuda", gen._device_index)
        self.dev = dev
        mv, P = gen.max_voxels, gen.max_num_points
        self.handle = _lib.Handle(gen.geometry, mv, P, n_feat, 1, cap, gen._device_index)
        with torch.cuda.device(dev):
            self.h_pts = torch.empty((cap, n_feat), dtype=torch.float32).pin_memory()
            self.h_offs = torch.zeros((2,), dtype=torch.int32).pin_memory()
            self.h_vnum = torch.zeros((1,), dtype=torch.int32).pin_memory()
            self.np_pts, self.np_offs, self.np_vnum = self.h_pts.numpy(), self.h_offs.numpy(), self.h_vnum.numpy()
            self.d_pts = torch.empty((cap, n_feat), dtype=torch.float32, device=dev)
            self.d_offs = torch.zeros((2,), dtype=torch.int32, device=dev)
            self.voxels = torch.empty((mv, P, n_feat), dtype=torch.float32, device=dev)
            self.coors = torch.empty((mv, 3), dtype=torch.int32, device=dev)
            self.num = torch.empty((mv,), dtype=torch.int32, device=dev)
            self.vnum = torch.zeros((1,), dtype=torch.int32, device=dev)
            self.event = torch.cuda.Event()
            self.copy_event, self.copy_pending = torch.cuda.Event(), False
            self.stream = torch.cuda.Stream(dev)      # capture stream; replays run on the caller's current stream
            self.graphs = {}                          # staged point rows (multiple of 4096) -> (CUDAGraph, cudaGraphExec_t)
            # raw handles for pf_frame_run / pf_frame_copy_out (one foreign call each per frame); a torch event owns its
            # cudaEvent_t from the first record on
            self.event.record(self.stream)
            self.copy_event.record(self.stream)
            self.stream.synchronize()
            self.ev_done, self.ev_copy = self.event.cuda_event, self.copy_event.cuda_event
            self.p_stage, self.p_offs, self.p_vnum = self.h_pts.data_ptr(), self.h_offs.data_ptr(), self.h_vnum.data_ptr()
            self.p_voxels, self.p_coors, self.p_num = self.voxels.data_ptr(), self.coors.data_ptr(), self.num.data_ptr()
            self.p_dpts = self.d_pts.data_ptr()
            self.c_v = C.c_int32(0)
            self.lib = self.handle.lib

    def _body(self, src_host: bool, rows: int):
        torch = _torch()
        h = self.handle
        if src_host:
            if rows:                                  # rows == 0: pf_frame_run uploads the points in front of the graph
                self.d_pts[:rows].copy_(self.h_pts[:rows], non_blocking=True)
            self.d_offs.copy_(self.h_offs, non_blocking=True)
        stream = torch.cuda.current_stream(self.dev).cuda_stream
        _lib.check(h.lib.pf_voxelize(h.ptr, self.d_pts.data_ptr(), self.d_offs.data_ptr(), 1, self.cap, self.F, 0, 3,
                                     self.voxels.data_ptr(), self.coors.data_ptr(), self.num.data_ptr(),
                                     self.vnum.data_ptr(), stream), "pf_voxelize")
        self.h_vnum.copy_(self.vnum, non_blocking=True)

    def graph_for(self, rows: int):
        """The graph whose H2D node moves `rows` staged point rows (rows is a multiple of 4096 <= cap); rows == 0: the graph
        without an upload of the points (offsets H2D, kernels, D2H of V), for the chunked upload of pf_frame_run."""
        torch = _torch()
        g = self.graphs.get(rows)
        if g is None:
            torch.cuda.synchronize(self.dev)          # earlier replays on the caller's stream use the same buffers
            with torch.cuda.device(self.dev), torch.cuda.stream(self.stream):
                self._body(True, rows)                # warm-up outside capture (first-use initialisation); idempotent
                self.stream.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=self.stream):
                    self._body(True, rows)
            torch.cuda.synchronize(self.dev)
            g = (g, g.raw_cuda_graph_exec())
            self.graphs[rows] = g
        return g

    def close(self):
        self.graphs.clear()
        self.handle.close()


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
        self._voxel_size_cfg = list(config["voxel_size"])
        # the reference writes these four keys back; AnchorAssigner, the datasets and the networks read them
        # (framework/voxel_generator.py:23-26)
        config["detection_range"] = geom.detection_range
        config["detection_offset"] = geom.offset
        config["detection_range_diff"] = geom.range_diff
        config["grid_size"] = geom.grid_size
        dev = device if device is not None else config.get("device", 0)
        self._device_index = self._parse_device(dev)
        self._cap_hint = int(max_points_per_frame) if max_points_per_frame else 0
        self._handles = {}      # F -> _lib.Handle of the batched / CUDA-input path
        self._paths = {}        # (F, capacity) -> _FramePath of the single-frame path
        self._pool = []         # _OutBlock
        self._pool_F = None
        self._fused = {}        # (F, capacity, id(pfn)) -> _FusedPath (generate_and_scatter)
        self._threads = _stage_threads()
        self._cuda_ok = False
        self._batch_stage = None   # pinned staging of generate_batch for lists of numpy frames
        import threading
        self._lock = threading.Lock()   # the reference's generate is a pure function and therefore thread-safe; this one owns
                                        # buffers, so concurrent calls on ONE instance are serialised (use one instance per thread
                                        # to run them side by side)
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

    def _require_cuda(self):
        if self._cuda_ok:
            return
        if not _torch().cuda.is_available():
            raise _lib.PillarFrontError("no CUDA device: the pillar front end has no CPU fallback")
        self._cuda_ok = True

    def _handle(self, n_frames: int, n_feat: int, max_pts: int) -> "_lib.Handle":
        torch = _torch()
        self._require_cuda()
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

    @staticmethod
    def _check_points(points):
        """Validate one frame; returns ("numpy" | "torch", N, F)."""
        torch = _torch()
        if isinstance(points, np.ndarray):
            if points.dtype != np.float32:
                raise TypeError(f"points must be float32, got {points.dtype}")
            if points.ndim != 2 or points.shape[1] < 3:
                raise ValueError(f"points must have shape (N, F>=3), got {points.shape}")
            kind = "numpy"
        elif isinstance(points, torch.Tensor):
            if points.dtype != torch.float32:
                raise TypeError(f"points must be float32, got {points.dtype}")
            if points.dim() != 2 or points.shape[1] < 3:
                raise ValueError(f"points must have shape (N, F>=3), got {tuple(points.shape)}")
            kind = "torch"
        else:
            raise TypeError(f"points must be a numpy array or a torch tensor, got {type(points).__name__}")
        n_feat = int(points.shape[1])
        if not 3 <= n_feat <= 16:
            raise ValueError(f"point features must be 3..16, got {n_feat}")
        return kind, int(points.shape[0]), n_feat

    def _to_device(self, points):
        """Validate one frame and return a contiguous fp32 CUDA tensor (N, F) (batched / CUDA-input path)."""
        torch = _torch()
        self._require_cuda()
        kind, _, _ = self._check_points(points)
        dev = torch.device("cuda", self._device_index)
        if kind == "numpy":
            t = torch.from_numpy(np.ascontiguousarray(points))
            return t.to(dev, non_blocking=True), False
        return points.contiguous().to(dev, non_blocking=True), points.is_cuda

    # ---- single-frame fast path ------------------------------------------------------------------------
    def _capacity(self, n: int) -> int:
        want = max(n, self._cap_hint, 1)
        return (want + _BUCKET - 1) // _BUCKET * _BUCKET

    def _path(self, n_feat: int, n: int) -> _FramePath:
        self._require_cuda()
        key = (n_feat, self._capacity(n))
        p = self._paths.get(key)
        if p is None:
            p = _FramePath(self, n_feat, key[1])
            self._paths[key] = p
        return p

    def _out_block(self) -> "_OutBlock | None":
        for b in self._pool:
            if b.free():
                return b
        if len(self._pool) < _POOL_MAX:
            b = _OutBlock(_torch(), self.max_voxels, self.max_num_points, self._pool_F)
            self._pool.append(b)
            return b
        return None

    def _generate_frame(self, points, n: int, n_feat: int, device_outputs: bool):
        """Stage + replay + wait (pf_frame_run), then hand out exactly V rows (pf_frame_copy_out): two foreign calls."""
        torch = _torch()
        path = self._path(n_feat, n)
        src = points if points.flags.c_contiguous else np.ascontiguousarray(points)
        # the upload of the staged frame is a node of the graph (rows = what it moves); PF_FRAME_UPLOAD=chunked: see above
        chunked = _CHUNKED_UPLOAD
        rows = 0 if chunked else min(path.cap, max(4096, (n + 4095) // 4096 * 4096))
        _, exec_ptr = path.graph_for(rows)
        dev_idx = self._device_index
        raw = _raw_stream(dev_idx)                    # replay on the caller's current stream, like every other call
        lib = path.lib
        # a hand-out copy of the previous call may still be running, possibly on another stream: the replay waits for it
        rc = lib.pf_frame_run(dev_idx, exec_ptr, raw, path.ev_done, path.ev_copy if path.copy_pending else None, path.p_stage,
                              src.ctypes.data if n else None, src.nbytes if n else 0, self._threads,
                              path.p_dpts if chunked else None, path.p_offs, n, path.p_vnum, C.byref(path.c_v))
        if rc:
            _lib.check(rc, "pf_frame_run")
        path.copy_pending = False
        v = path.c_v.value
        P = self.max_num_points
        if device_outputs:                            # fresh tensors of exactly V rows: nothing aliases the workspace
            dev = path.dev
            out = (torch.empty((v, P, n_feat), dtype=torch.float32, device=dev),
                   torch.empty((v, 3), dtype=torch.int32, device=dev), torch.empty((v,), dtype=torch.int32, device=dev))
            rc = lib.pf_frame_copy_out(dev_idx, out[0].data_ptr(), path.p_voxels, v * P * n_feat * 4, out[1].data_ptr(),
                                       path.p_coors, v * 12, out[2].data_ptr(), path.p_num, v * 4, raw, path.ev_copy, 0)
            if rc:
                _lib.check(rc, "pf_frame_copy_out")
            path.copy_pending = True
            return out, v
        if self._pool and self._pool_F != n_feat:
            self._pool = []
        self._pool_F = n_feat
        blk = self._out_block()
        if blk is not None:                           # D2H of exactly V rows, straight into the arrays handed out
            rc = lib.pf_frame_copy_out(dev_idx, blk.p_v, path.p_voxels, v * P * n_feat * 4, blk.p_c, path.p_coors, v * 12,
                                       blk.p_n, path.p_num, v * 4, raw, None, 1)
            if rc:
                _lib.check(rc, "pf_frame_copy_out")
            return (blk.nv[:v], blk.nc[:v], blk.nn[:v]), v
        torch.cuda.current_stream(path.dev).synchronize()
        return (path.voxels[:v].cpu().numpy(), path.coors[:v].cpu().numpy(), path.num[:v].cpu().numpy()), v

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
        with torch.cuda.device(dev):
            pts = frames_dev[0] if n_frames == 1 else torch.cat(frames_dev, dim=0)
            if pts.numel() == 0:
                pts = torch.zeros((1, n_feat), dtype=torch.float32, device=dev)
            offs = torch.tensor(np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32), device=dev)
        return self._voxelize_batch(pts, offs, n_frames, n_feat, max(sizes + [1]), compact, coors_stride)

    def _run_host(self, frames, n_feat: int, compact: bool, coors_stride: int):
        """A list of C-contiguous float32 numpy frames: ONE pf_upload_frames call stages them (streaming stores, a few host
        threads) into a pinned buffer and enqueues one DMA per frame, instead of a pageable copy + wait per frame."""
        torch = _torch()
        dev = torch.device("cuda", self._device_index)
        n_frames = len(frames)
        sizes = [int(f.shape[0]) for f in frames]
        total = max(1, sum(sizes))
        st = self._batch_stage
        if st is None or st["F"] != n_feat or st["rows"] < total or st["frames"] < n_frames:
            rows = max(total, st["rows"] if st and st["F"] == n_feat else 0)
            rows = (rows + _BUCKET - 1) // _BUCKET * _BUCKET
            slots = max(n_frames, st["frames"] if st else 0)
            torch.cuda.synchronize(dev)                       # the DMA of an earlier batch may still read the old buffer
            with torch.cuda.device(dev):
                st = dict(F=n_feat, rows=rows, frames=slots, h_pts=torch.empty((rows, n_feat), dtype=torch.float32).pin_memory(),
                          h_offs=torch.zeros((slots + 1,), dtype=torch.int32).pin_memory(), event=torch.cuda.Event())
            st["ptrs"], st["nrows"], st["kinds"] = (C.c_void_p * slots)(), (C.c_int64 * slots)(), (C.c_int32 * slots)()
            st["busy"] = False
            self._batch_stage = st
        if st["busy"]:
            st["event"].synchronize()                         # the previous batch's DMA out of the staging buffer
        for b, f in enumerate(frames):
            st["ptrs"][b], st["nrows"][b], st["kinds"][b] = f.ctypes.data if sizes[b] else None, sizes[b], 0
        with torch.cuda.device(dev):
            pts = torch.empty((total, n_feat), dtype=torch.float32, device=dev)
            raw = _raw_stream(self._device_index)
            lib = _lib.load()
            _lib.check(lib.pf_upload_frames(pts.data_ptr(), st["h_pts"].data_ptr(), st["ptrs"], st["nrows"], st["kinds"], n_frames,
                                            n_frames, n_feat * 4, self._threads, st["h_offs"].data_ptr(), raw), "pf_upload_frames")
            offs = st["h_offs"][:n_frames + 1].to(dev, non_blocking=True)
            st["event"].record(torch.cuda.current_stream(dev))
            st["busy"] = True
        return self._voxelize_batch(pts, offs, n_frames, n_feat, max(sizes + [1]), compact, coors_stride)

    def _voxelize_batch(self, pts, offs, n_frames: int, n_feat: int, max_pts: int, compact: bool, coors_stride: int):
        torch = _torch()
        dev = torch.device("cuda", self._device_index)
        h = self._handle(n_frames, n_feat, max_pts)
        with torch.cuda.device(dev):
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
        t0 = time.time()
        self._require_cuda()
        kind, n, n_feat = self._check_points(points)
        on_dev = kind == "torch" and points.is_cuda
        if device_outputs is None:
            device_outputs = on_dev
        if kind == "torch" and not on_dev:
            points, kind = points.contiguous().numpy(), "numpy"
        if kind == "numpy":
            with self._lock:
                (voxels, coors, num), v = self._generate_frame(points, n, n_feat, bool(device_outputs))
        else:                                          # CUDA tensor in: no staging, eager launch on the caller's stream
            dv, dc, dn, vnum = self._run([points.contiguous().to(_torch().device("cuda", self._device_index))],
                                         compact=False, coors_stride=3)
            v = int(vnum.item())
            voxels, coors, num = dv[:v], dc[:v], dn[:v]
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
        stride = 4 if len(frames) > 1 else 3       # merge_second_batch pads the frame id only when B > 1
        kinds = [self._check_points(f) for f in frames]
        if all(k[0] == "numpy" for k in kinds) and len({k[2] for k in kinds}) == 1:
            self._require_cuda()
            host = [f if f.flags.c_contiguous else np.ascontiguousarray(f) for f in frames]
            with self._lock:
                voxels, coors, num, vnum = self._run_host(host, kinds[0][2], compact=True, coors_stride=stride)
        else:
            devs = [self._to_device(f)[0] for f in frames]
            voxels, coors, num, vnum = self._run(devs, compact=True, coors_stride=stride)
        counts = vnum.cpu().numpy()
        total = int(counts.sum())
        out = {"voxels": voxels[:total], "coordinates": coors[:total], "num_points_per_voxel": num[:total],
               "voxel_num": counts}
        if not device_outputs:
            out = {k: (v.cpu().numpy() if hasattr(v, "cpu") else v) for k, v in out.items()}
        return out

    # ---- one graph per frame: voxelise + decorate + PFN + canvas --------------------------------------
    def generate_and_scatter(self, points, pfn, n_channels=64):
        """Single-frame fast path for inference loops (ros_node.py:57-75, train.py infer): numpy frame in, BEV canvas
        ``(1, C, nx, ny)`` on the device out, as ONE CUDA graph (H2D, voxelise, decorate + fused PFN, canvas fill, D2H of
        the voxel count) and one synchronisation.  ``pfn`` is a :class:`PointNet` in eval mode; its folded weights are
        read when the graph is first captured (call again after changing them: the cache key follows their versions).
        Returns ``(canvas, coors (V, 4) int32 CUDA, voxel_num)``; the tensors are views of per-instance buffers, valid
        until the next call."""
        from .frame_path import FusedFramePath
        self._require_cuda()
        kind, n, n_feat = self._check_points(points)
        if kind != "numpy":
            points = points.detach().cpu().contiguous().numpy()
        if n_feat != 4:
            raise ValueError("generate_and_scatter needs 4 floats per point (the fused PFN kernel)")
        fold = pfn.folded()
        key = (self._capacity(n), id(pfn), tuple(t.data_ptr() for t in fold), getattr(pfn, "_fold_key", None))
        fp = self._fused.get(key)
        if fp is None:
            for old in [k for k in self._fused if k[0] == key[0] and k[1] == key[1]]:
                self._fused.pop(old).close()
            fp = FusedFramePath(self, key[0], fold, n_channels)
            self._fused[key] = fp
        with self._lock:
            return fp.run(points, n, self._threads)

    def close(self):
        for h in self._handles.values():
            h.close()
        self._handles.clear()
        for p in self._paths.values():
            p.close()
        self._paths.clear()
        for p in self._fused.values():
            p.close()
        self._fused.clear()
        self._pool = []


class VoxelGenerator_trt(VoxelGenerator):
    """framework/voxel_generator.py:43-79: same, but generate() also returns voxel_num."""
    _returns_voxel_num = True
