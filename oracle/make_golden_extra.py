"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Extra golden vectors from the REFERENCE ITSELF for geometries no shipped
config uses but the reference code supports: a true 3-D grid (nz > 1) and point rows with F != 4.

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_extra        (build container only: needs /root/reference)

Writes tests/golden/extra_<case>.npz (inputs + the reference's VoxelGenerator.generate outputs).
"""
from __future__ import annotations

import importlib

import numpy as np

from oracle.make_golden import GOLD, ref_modules

cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")

CASES = [
    # name, detection_range, voxel_size, max_voxels, P, N, F, seed
    ("3d_overflow", [-8, -6, -2, 8, 6, 2], [0.25, 0.2, 0.5], 600, 5, 6000, 4, 901),
    ("3d_fits", [-8, -6, -2, 8, 6, 2], [0.5, 0.5, 1.0], 4000, 8, 3000, 4, 902),
    ("3d_f6", [-8, -6, -2, 8, 6, 2], [0.25, 0.2, 0.5], 900, 3, 2500, 6, 903),
    ("2d_f3", [-8, -6, -2, 8, 6, 2], [0.25, 0.2, 4.0], 700, 4, 2500, 3, 904),
]


def main():
    VoxelGenerator, _, _ = ref_modules()
    for name, rng_, vs, mv, P, n, F, seed in CASES:
        cfg = dict(detection_range=list(rng_), voxel_size=list(vs), max_voxels=mv, max_num_points=P,
                   num_point_features=F, batch_size=1, device=0)
        geom = cfgmod.derive_grid(cfg)
        base = synth.s3_uniform(geom, n, seed)
        rng = np.random.default_rng(seed + 1)
        if F == 3:
            pts = base[:, :3]
        elif F > 4:
            pts = np.concatenate([base, rng.random((n, F - 4), dtype=np.float32)], axis=1)
        else:
            pts = base
        pts = np.ascontiguousarray(pts, np.float32)
        vg = VoxelGenerator(cfg)
        voxels, coors, num = vg.generate(pts)
        np.savez_compressed(GOLD / f"extra_{name}.npz", points=pts, voxels=voxels, coors=coors, num=num,
                            detection_range=np.asarray(rng_, np.float32), voxel_size_cfg=np.asarray(vs, np.float64),
                            max_voxels=np.int32(mv), max_points=np.int32(P), grid_size=vg.grid_size, offset=vg.offset)
        print(f"extra {name:12s} grid={vg.grid_size.tolist()} N={n} F={F} V={len(num)} kept={int(num.sum())}")


if __name__ == "__main__":
    main()
