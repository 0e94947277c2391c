"""B200-native point-cloud front end (voxelise -> pillar decoration -> BEV scatter) for PointPillars.

Drop-in for the reference's ``framework/voxel_generator.py`` and the decoration / ``PointPillarsScatter``
parts of ``networks/pointpillars8_*.py``; every compute call goes through the C-ABI CUDA library
``libpillarfront.so`` (include/pillarfront.h).  There is no CPU fallback.

The directory name is not a Python identifier; import it as ``pillarfront`` (repo-root shim) or with
``importlib.import_module("3d_object_detection_b200")``.
"""
from .config import GRIDS, CONFIG_GRID, GridGeometry, builtin_config, derive_grid, load_config  # noqa: F401

__all__ = ["GRIDS", "CONFIG_GRID", "GridGeometry", "builtin_config", "derive_grid", "load_config"]
