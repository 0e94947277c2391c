"""ctypes binding of libpillarfront.so (C ABI: include/pillarfront.h).

There is no fallback: if the CUDA library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "lib" / "libpillarfront.so"

EXPORTS = ["pf_version", "pf_last_error_string", "pf_create", "pf_destroy", "pf_workspace_bytes",
           "pf_voxelize", "pf_voxelize_decorate", "pf_voxelize_pfn", "pf_pfn", "pf_decorate", "pf_scatter", "pf_scatter_static", "pf_scatter_from_map",
           "pf_scatter_backward", "pf_anchor_mask", "pf_nms", "pf_nms_scratch_bytes",
           "pf_debug_stage_timing", "pf_debug_stage_times", "pf_debug_set_fill_variant", "pf_debug_set_pdl", "pf_discard_cell_map", "pf_stage_copy", "pf_copy3", "pf_frame_run", "pf_frame_copy_out", "pf_upload_frames", "pf_runner_create", "pf_runner_set_slot", "pf_runner_submit",
           "pf_runner_wait", "pf_runner_destroy", "pf_runner_last_error",
           "pf_postprocess", "pf_post_scratch_bytes", "pf_post_scratch_layout"]


class GridDesc(C.Structure):
    """pf_grid_desc"""
    _fields_ = [("offset", C.c_float * 3), ("voxel_size", C.c_float * 3), ("grid", C.c_int32 * 3),
                ("max_voxels", C.c_int32), ("max_points", C.c_int32), ("num_features", C.c_int32)]


class PostDesc(C.Structure):
    """pf_post_desc"""
    _fields_ = [("n_classes", C.c_int32), ("class_range", (C.c_int32 * 2) * 8), ("n_cls_channels", C.c_int32),
                ("pre_max", C.c_int32), ("post_max", C.c_int32), ("score_threshold", C.c_float),
                ("iou_threshold", C.c_float), ("center_limit", C.c_double * 6)]


# pf_post_region: name -> (index, torch dtype of the region, columns per row)
POST_REGION_NAMES = ["cand", "cand_count", "sel_key", "sel_num", "seg_offsets", "sel_anchor", "sel_score", "sel_box",
                     "sel_standup", "sel_dir", "sel_rank", "keep", "keep_num", "nms_scratch"]
_POST_TYPES = {"cand": ("int64", 1), "cand_count": ("int32", 1), "sel_key": ("int64", 1), "sel_num": ("int32", 1),
               "seg_offsets": ("int32", 1), "sel_anchor": ("int32", 1), "sel_score": ("float32", 1),
               "sel_box": ("float32", 7), "sel_standup": ("float32", 4), "sel_dir": ("int32", 1),
               "sel_rank": ("float32", 1), "keep": ("int32", 1), "keep_num": ("int32", 1), "nms_scratch": ("uint8", 1)}
POST_REGIONS = {n: (i, *_POST_TYPES[n]) for i, n in enumerate(POST_REGION_NAMES)}


class PillarFrontError(RuntimeError):
    pass


_lib = None
_p = C.c_void_p


def load():
    """Load the shared library once; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise PillarFrontError(
            f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
            "(python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback.")
    L = C.CDLL(str(LIB_PATH))
    L.pf_version.restype = C.c_int
    L.pf_version.argtypes = []
    L.pf_last_error_string.restype = C.c_char_p
    L.pf_last_error_string.argtypes = []
    L.pf_create.restype = C.c_int
    L.pf_create.argtypes = [C.POINTER(GridDesc), C.c_int, C.c_int, C.c_int, C.POINTER(_p)]
    L.pf_destroy.restype = C.c_int
    L.pf_destroy.argtypes = [_p]
    L.pf_workspace_bytes.restype = C.c_size_t
    L.pf_workspace_bytes.argtypes = [_p]
    L.pf_voxelize.restype = C.c_int
    L.pf_voxelize.argtypes = [_p, _p, _p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _p, _p, _p, _p, _p]
    L.pf_voxelize_decorate.restype = C.c_int
    L.pf_voxelize_decorate.argtypes = [_p, _p, _p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_float, C.c_float, C.c_float, C.c_float, C.c_int,
                                       _p, _p, _p, _p, _p, _p]
    L.pf_voxelize_pfn.restype = C.c_int
    L.pf_voxelize_pfn.argtypes = [_p, _p, _p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                  C.c_float, C.c_float, C.c_float, C.c_float, C.c_int,
                                  _p, _p, _p, C.c_int, _p, _p, _p, _p, _p, _p, _p]
    L.pf_pfn.restype = C.c_int
    L.pf_pfn.argtypes = [_p, _p, _p, C.c_int, C.c_int64, C.c_int, C.c_int,
                         C.c_float, C.c_float, C.c_float, C.c_float, _p, _p, _p, C.c_int, _p, _p]
    L.pf_decorate.restype = C.c_int
    L.pf_decorate.argtypes = [_p, _p, _p, C.c_int, C.c_int64, C.c_int, C.c_int,
                              C.c_float, C.c_float, C.c_float, C.c_float, _p, _p]
    L.pf_scatter.restype = C.c_int
    L.pf_scatter.argtypes = [_p, _p, C.c_int64, C.c_int64, _p, C.c_int, C.c_int64, C.c_int, C.c_int, _p, _p]
    L.pf_scatter_static.restype = C.c_int
    L.pf_scatter_static.argtypes = [_p, C.c_int64, C.c_int64, _p, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p, _p]
    L.pf_scatter_from_map.restype = C.c_int
    L.pf_scatter_from_map.argtypes = [_p, _p, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int, _p, _p]
    L.pf_scatter_backward.restype = C.c_int
    L.pf_scatter_backward.argtypes = [_p, _p, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p,
                                      C.c_int64, C.c_int64, _p]
    L.pf_anchor_mask.restype = C.c_int
    L.pf_anchor_mask.argtypes = [_p, C.c_int, _p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, _p, C.c_int64, _p, _p, _p]
    L.pf_debug_stage_timing.restype = C.c_int
    L.pf_debug_stage_timing.argtypes = [_p, C.c_int]
    L.pf_debug_stage_times.restype = C.c_int
    L.pf_debug_stage_times.argtypes = [_p, C.POINTER(C.c_float), C.c_int]
    L.pf_stage_copy.restype = C.c_int
    L.pf_stage_copy.argtypes = [_p, _p, C.c_size_t, C.c_int]
    L.pf_frame_run.restype = C.c_int
    L.pf_frame_run.argtypes = [C.c_int, _p, _p, _p, _p, _p, _p, C.c_size_t, C.c_int, _p, _p, C.c_int, _p, C.POINTER(C.c_int32)]
    L.pf_frame_copy_out.restype = C.c_int
    L.pf_frame_copy_out.argtypes = [C.c_int, _p, _p, C.c_size_t, _p, _p, C.c_size_t, _p, _p, C.c_size_t, _p, _p, C.c_int]
    L.pf_upload_frames.restype = C.c_int
    L.pf_upload_frames.argtypes = [_p, _p, C.POINTER(_p), C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int,
                                   C.c_int, _p, _p]
    L.pf_runner_create.restype = C.c_int
    L.pf_runner_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_p)]
    L.pf_runner_set_slot.restype = C.c_int
    L.pf_runner_set_slot.argtypes = [_p, C.c_int, _p, _p, _p, _p, _p]
    L.pf_runner_submit.restype = C.c_longlong
    L.pf_runner_submit.argtypes = [_p, C.POINTER(_p), C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.c_int]
    L.pf_runner_wait.restype = C.c_int
    L.pf_runner_wait.argtypes = [_p, C.c_longlong, C.POINTER(C.c_float)]
    L.pf_runner_destroy.restype = C.c_int
    L.pf_runner_destroy.argtypes = [_p]
    L.pf_runner_last_error.restype = C.c_char_p
    L.pf_runner_last_error.argtypes = []
    L.pf_copy3.restype = C.c_int
    L.pf_copy3.argtypes = [_p, _p, C.c_size_t, _p, _p, C.c_size_t, _p, _p, C.c_size_t, _p]
    L.pf_debug_set_fill_variant.restype = C.c_int
    L.pf_debug_set_fill_variant.argtypes = [C.c_int]
    L.pf_debug_set_pdl.restype = C.c_int
    L.pf_debug_set_pdl.argtypes = [C.c_int]
    L.pf_discard_cell_map.restype = C.c_int
    L.pf_discard_cell_map.argtypes = [_p, _p]
    L.pf_nms_scratch_bytes.restype = C.c_size_t
    L.pf_nms_scratch_bytes.argtypes = [C.c_int, C.c_int]
    L.pf_nms.restype = C.c_int
    L.pf_nms.argtypes = [_p, _p, _p, C.c_int, C.c_int64, C.c_int, C.c_float, C.c_int, _p, _p, _p, _p]
    L.pf_post_scratch_bytes.restype = C.c_size_t
    L.pf_post_scratch_bytes.argtypes = [C.POINTER(PostDesc), C.c_int, C.c_int64]
    L.pf_post_scratch_layout.restype = C.c_size_t
    L.pf_post_scratch_layout.argtypes = [C.POINTER(PostDesc), C.c_int, C.c_int64, C.POINTER(C.c_size_t)]
    L.pf_postprocess.restype = C.c_int
    L.pf_postprocess.argtypes = [C.POINTER(PostDesc), _p, _p, _p, _p, _p, C.c_int, C.c_int64, _p, _p, _p, _p, _p, _p, _p]
    _lib = L
    return L


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().pf_last_error_string().decode("utf-8", "replace")
        raise PillarFrontError(f"{what or 'libpillarfront'} failed (status {rc}): {msg}")


def make_post_desc(class_ranges, n_cls_channels, pre_max, post_max, score_threshold, iou_threshold, center_limit) -> PostDesc:
    if not 1 <= len(class_ranges) <= 8:
        raise ValueError(f"{len(class_ranges)} classes: pf_postprocess handles 1..8")
    d = PostDesc()
    d.n_classes = len(class_ranges)
    for c, (lo, hi) in enumerate(class_ranges):
        d.class_range[c][0], d.class_range[c][1] = int(lo), int(hi)
    d.n_cls_channels, d.pre_max, d.post_max = int(n_cls_channels), int(pre_max), int(post_max)
    d.score_threshold, d.iou_threshold = float(score_threshold), float(iou_threshold)
    for i in range(6):
        d.center_limit[i] = float(center_limit[i])
    return d


def make_desc(geom, max_voxels: int, max_points: int, num_features: int) -> GridDesc:
    d = GridDesc()
    for k in range(3):
        d.offset[k] = float(geom.offset[k])
        d.voxel_size[k] = float(geom.voxel_size[k])
        d.grid[k] = int(geom.grid_size[k])
    d.max_voxels, d.max_points, d.num_features = int(max_voxels), int(max_points), int(num_features)
    return d


class Handle:
    """Owns one pf_handle (workspace bound to one device; use one per stream)."""

    def __init__(self, geom, max_voxels, max_points, num_features, max_frames, max_points_per_frame, device):
        self.lib = load()
        self.desc = make_desc(geom, max_voxels, max_points, num_features)
        self.max_frames = int(max_frames)
        self.max_points_per_frame = int(max_points_per_frame)
        self.device = int(device)
        self._h = _p()
        check(self.lib.pf_create(C.byref(self.desc), self.max_frames, self.max_points_per_frame,
                                 self.device, C.byref(self._h)), "pf_create")

    @property
    def ptr(self):
        return self._h

    @property
    def workspace_bytes(self) -> int:
        return int(self.lib.pf_workspace_bytes(self._h))

    def close(self):
        if self._h:
            self.lib.pf_destroy(self._h)
            self._h = _p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- cheap torch plumbing for the per-call host paths (a Stream object or a device context costs 5-8 us each) ----------
_RAW_STREAM = None


def raw_stream(device_index: int) -> int:
    """cudaStream_t of torch's current stream on the device, without building a torch.cuda.Stream object."""
    global _RAW_STREAM
    if _RAW_STREAM is None:
        import torch
        fast = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        _RAW_STREAM = fast if fast is not None else (lambda idx: torch.cuda.current_stream(idx).cuda_stream)
    if device_index is None:
        import torch
        device_index = torch.cuda.current_device()
    return _RAW_STREAM(device_index)


class _NoCtx:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


_NO_CTX = _NoCtx()


def on_device(device):
    """Context that makes `device` current for handle-free library calls; free when it already is."""
    import torch
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if torch.cuda.current_device() == idx:
        return _NO_CTX
    return torch.cuda.device(idx)
