"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Golden vectors for the anchor occupancy mask ("next" row N2) made by the
REFERENCE ITSELF: anchors from AnchorAssigner.generate + rbbox2d_to_near_bbox, anchor rectangles from
get_anchor_coor, and the mask from the CPU branch of AnchorAssigner.create_mask
(framework/anchor_assigner.py:322-335: sparse_sum_for_anchors_mask -> cumsum(0) -> cumsum(1) ->
fused_get_anchors_area > 0), which is the same arithmetic as the numba-CUDA branch (box_np_ops.py:168-257) that
cannot run in the build container.

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_anchor        (build container only: needs /root/reference)

Writes tests/golden/anchor_<case>.npz.
"""
from __future__ import annotations

import importlib

import numpy as np

from oracle.make_golden import GOLD, ref_modules

cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")

CLASSES = {   # sizes / rotations of framework/anchor_assigner.py:222-246, one feature map per case
    "vehicle": ([[4.6, 2.10, 1.8], [7.5, 2.6, 2.9], [12.6, 2.9, 3.8]], [0, 1.5707963267948966]),
    "pedestrian": ([[0.96874749, 0.9645992, 1.81212425]], [0]),
    "cyclist": ([[2.02032733, 0.98075615, 1.72027404]], [0, 1.5707963267948966]),
}

CASES = [
    # name, detection_range, voxel_size, max_voxels, P, N points, feature map, seed, generator
    ("pillar_small", [-16.0, -12.8, -2.5, 16.0, 12.8, 8.5], [0.2, 0.2, 11], 3000, 15, 9000, [80, 64, 1], 1201, "s1"),
    ("pillar_sparse", [-16.0, -12.8, -2.5, 16.0, 12.8, 8.5], [0.2, 0.2, 11], 3000, 15, 300, [80, 64, 1], 1202, "s3"),
    ("grid3d", [-8, -6, -2, 8, 6, 2], [0.25, 0.2, 0.5], 2500, 5, 6000, [32, 30, 1], 1203, "s3"),
]


def ref_anchors(assigner_cls, box_np_ops, cfg, fmap):
    """The anchor loop of AnchorAssigner.__init__ (framework/anchor_assigner.py:252-289) driven through the
    reference's own generate(); __init__ itself cannot run here because it calls cuda.to_device."""
    a = object.__new__(assigner_cls)
    a.anchor_offsets = cfg["detection_offset"]
    bvs = []
    real_meshgrid = np.meshgrid      # numpy >= 2 returns a tuple; generate() (:313,318) does `locations + [...]`
    np.meshgrid = lambda *xs, **kw: list(real_meshgrid(*xs, **kw))
    for cls, (sizes, rotations) in CLASSES.items():
        lst = []
        for size in sizes:
            fm = np.array(fmap, dtype=np.float32)
            strides = cfg["detection_range_diff"] / fm
            for rot in rotations:
                lst.append(a.generate(size, rot, fm, strides).reshape([-1, 7]))
        lst = np.concatenate(lst)
        bvs.append(box_np_ops.rbbox2d_to_near_bbox(lst[:, [0, 1, 3, 4, 6]]))
    np.meshgrid = real_meshgrid
    return np.concatenate(bvs)


def main():
    VoxelGenerator, _, _ = ref_modules()
    import framework.anchor_assigner as ra
    import framework.box_np_ops as box_np_ops
    for name, rng_, vs, mv, P, n, fmap, seed, gen in CASES:
        cfg = dict(detection_range=list(rng_), voxel_size=list(vs), max_voxels=mv, max_num_points=P,
                   num_point_features=4, batch_size=1, device=0)
        geom = cfgmod.derive_grid(cfg)
        pts = synth.s2_ring(geom, n, seed) if gen == "s1" else synth.s3_uniform(geom, n, seed)
        vg = VoxelGenerator(cfg)      # writes detection_offset / detection_range_diff / grid_size into cfg
        _, coors, _ = vg.generate(np.ascontiguousarray(pts, np.float32))
        anchors_bv = ref_anchors(ra.AnchorAssigner, box_np_ops, cfg, fmap)
        voxel_size = np.array(cfg["voxel_size"], dtype=np.float32)
        offset = np.array(cfg["detection_offset"], dtype=np.float32)
        grid_size = vg.grid_size
        anchor_coors = box_np_ops.get_anchor_coor(anchors_bv, voxel_size, offset, grid_size)
        dense = box_np_ops.sparse_sum_for_anchors_mask(coors, tuple(grid_size[:-1]))
        dense = dense.cumsum(0)
        dense = dense.cumsum(1)
        area = box_np_ops.fused_get_anchors_area(dense, anchors_bv, voxel_size, offset, grid_size)
        mask = area > 0
        np.savez_compressed(GOLD / f"anchor_{name}.npz", coors=coors, anchors_bv=anchors_bv, anchor_coors=anchor_coors,
                            mask=mask, area=area, voxel_size=voxel_size, offset=offset, grid_size=grid_size)
        print(f"anchor {name:14s} grid={grid_size.tolist()} V={len(coors)} A={len(mask)} set={int(mask.sum())} "
              f"bv dtype={anchors_bv.dtype}")


if __name__ == "__main__":
    main()
