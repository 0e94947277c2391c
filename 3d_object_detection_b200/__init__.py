"""B200-native point-cloud front end (voxelise -> pillar decoration -> BEV scatter) for PointPillars.

Drop-in for the reference's ``framework/voxel_generator.py`` and the decoration / ``PointPillarsScatter``
parts of ``networks/pointpillars8_*.py``; every compute call goes through the C-ABI CUDA library
``libpillarfront.so`` (include/pillarfront.h).  There is no CPU fallback.

The directory name is not a Python identifier; import it as ``pillarfront`` (repo-root shim) or with
``importlib.import_module("3d_object_detection_b200")``.
"""
from .config import GRIDS, CONFIG_GRID, GridGeometry, builtin_config, derive_grid, load_config  # noqa: F401

_LAZY = {
    "VoxelGenerator": "voxel_generator", "VoxelGenerator_trt": "voxel_generator",
    "decorate": "pillar_ops", "decorate_constants": "pillar_ops", "PointNet": "pillar_ops",
    "PointPillarsScatter": "pillar_ops", "Scatter": "pillar_ops", "Scatter_cuda": "pillar_ops",
    "AnchorMask": "anchor_mask", "get_anchor_coor": "anchor_mask", "fused_get_anchors_mask_gpu": "anchor_mask",
    "merge_second_batch": "collate", "example_convert_to_torch": "collate",
    "nms_gpu": "nms", "nms": "nms", "nms_device": "nms",
    "Inference": "inference", "PostProcessor": "inference",
    "PillarFrontEnd": "frontend", "shard_frames": "sharding", "ShardedFrontEnd": "sharding", "place_devices": "sharding",
    "PillarFrontError": "_lib",
}

__all__ = ["GRIDS", "CONFIG_GRID", "GridGeometry", "builtin_config", "derive_grid", "load_config", *_LAZY]


def __getattr__(name):
    # torch-dependent modules load on first use so config/synth helpers stay importable anywhere
    if name in _LAZY:
        import importlib
        mod = importlib.import_module(f"{__name__}.{_LAZY[name]}")
        val = getattr(mod, name)
        globals()[name] = val
        return val
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
