"""``import pillarfront`` -- importable alias of the package directory ``3d_object_detection_b200/``
(whose name is not a Python identifier).  ``from pillarfront import VoxelGenerator`` is the one-line swap
for ``from framework.voxel_generator import VoxelGenerator`` in the reference's scripts."""
import importlib
import sys

_pkg = importlib.import_module("3d_object_detection_b200")
sys.modules[__name__] = _pkg
