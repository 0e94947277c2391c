"""Config loading and grid geometry for the pillar front end.

The reference keeps its settings in a flat JSON dict that constructors read *and mutate*
(SURVEY.md F8).  This module holds the two pieces of host logic every entry point shares:

* :func:`load_config` -- ``json.load`` that also accepts ``configs/nuscene.json``, which ships with a
  trailing comma.
* :func:`derive_grid` -- the grid geometry of ``VoxelGenerator.__init__``
  (reference ``framework/voxel_generator.py:7-15``), evaluated with numpy fp32 so that e.g. the nuScenes
  offset comes out as ``-38.399998`` exactly like the reference.

The built-in table :data:`GRIDS` carries the voxel parameters of every ``configs/*.json`` geometry so the
benchmark and tests do not need ``/root/reference`` at run time (it does not exist on the GPU box).
"""
from __future__ import annotations

import json
import re
from dataclasses import dataclass

import numpy as np

# Voxel parameters of the reference's shipped configs (configs/*.json; values read by
# framework/voxel_generator.py:7-22).  "frame_points" is the synthetic frame size BASELINE.json quotes.
GRIDS = {
    # ntusl_20cm / seven_20cm / eight_20cm / viewer / *_20cm_mly
    "20cm": dict(detection_range=[-80.0, -80.0, -2.5, 80.0, 80.0, 8.5], voxel_size=[0.2, 0.2, 11],
                 max_voxels=16000, max_num_points=15, num_point_features=4, frame_points=120000),
    # ntusl_10cm / ntusl_10cm_mly
    "10cm": dict(detection_range=[-80.0, -80.0, -2.5, 80.0, 80.0, 8.5], voxel_size=[0.1, 0.1, 11],
                 max_voxels=20000, max_num_points=10, num_point_features=4, frame_points=300000),
    # inhouse
    "16cm": dict(detection_range=[-80.0, -80.0, -2.5, 80.0, 80.0, 8.5], voxel_size=[0.16, 0.16, 11],
                 max_voxels=16000, max_num_points=15, num_point_features=4, frame_points=120000),
    # nuscene / config_nuscene
    "nuscene": dict(detection_range=[-30, -38.4, -3.5, 51.92, 38.4, 2.5], voxel_size=[0.16, 0.16, 6],
                    max_voxels=12000, max_num_points=100, num_point_features=4, frame_points=350000),
}

# config file stem -> GRIDS key
CONFIG_GRID = {
    "ntusl_20cm": "20cm", "ntusl_20cm_mly": "20cm", "seven_20cm": "20cm", "seven_20cm_mly": "20cm",
    "eight_20cm": "20cm", "eight_20cm_mly": "20cm", "viewer": "20cm",
    "ntusl_10cm": "10cm", "ntusl_10cm_mly": "10cm", "inhouse": "16cm",
    "nuscene": "nuscene", "config_nuscene": "nuscene",
}


def builtin_config(name: str) -> dict:
    """A fresh, mutable config dict for a grid key (``"20cm"``) or a config stem (``"seven_20cm"``)."""
    key = CONFIG_GRID.get(name, name)
    if key not in GRIDS:
        raise KeyError(f"unknown grid/config {name!r}; known: {sorted(GRIDS)} / {sorted(CONFIG_GRID)}")
    cfg = {k: (list(v) if isinstance(v, list) else v) for k, v in GRIDS[key].items()}
    cfg.update(batch_size=1, device=0)
    return cfg


def load_config(path) -> dict:
    """``json.load`` with tolerance for a trailing comma before ``}`` / ``]`` (configs/nuscene.json:24-25)."""
    with open(path, "r") as fh:
        text = fh.read()
    try:
        return json.loads(text)
    except json.JSONDecodeError:
        return json.loads(re.sub(r",(\s*[}\]])", r"\1", text))


@dataclass(frozen=True)
class GridGeometry:
    voxel_size: np.ndarray       # f32[3]
    detection_range: np.ndarray  # f32[6], snapped to an integer number of cells
    offset: np.ndarray           # f32[3], lower corner of cell (0,0,0)
    range_diff: np.ndarray       # f32[3], grid_size * voxel_size
    grid_size: np.ndarray        # i32[3], (nx, ny, nz)

    @property
    def cells(self) -> int:
        return int(np.prod(self.grid_size.astype(np.int64)))


def derive_grid(config: dict) -> GridGeometry:
    """Grid geometry exactly as the reference derives it (framework/voxel_generator.py:7-15).

    All arithmetic is numpy float32: extent / voxel_size is an fp32 divide truncated toward zero, the
    range is re-centred on its midpoint after truncation.  Nothing is written to ``config`` here; the
    ``VoxelGenerator`` constructor does that.
    """
    lim = np.asarray(config["detection_range"], dtype=np.float32)
    if lim.shape != (6,):
        raise ValueError("config['detection_range'] must hold 6 numbers (xyz-min, xyz-max)")
    vs = np.asarray(config["voxel_size"], dtype=np.float32)
    if vs.shape != (3,) or not np.all(vs > 0):
        raise ValueError("config['voxel_size'] must hold 3 positive numbers")
    lo, hi = lim[:3], lim[3:]
    mid = (hi + lo) / 2
    cells = ((hi - lo) / vs).astype(np.int32)
    if not np.all(cells > 0):
        raise ValueError(f"detection_range/voxel_size give an empty grid {cells}")
    span = cells.astype(np.float32) * vs
    origin = mid - span / 2
    snapped = np.concatenate((origin, origin + span), axis=0)
    return GridGeometry(voxel_size=vs, detection_range=snapped, offset=origin, range_diff=span,
                        grid_size=cells)
