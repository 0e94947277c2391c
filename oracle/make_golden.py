"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Generate tests/golden/ by running the REFERENCE ITSELF.

Run in the build container only (needs /root/reference, numba, torch-CPU):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden

It imports the reference's own ``framework.voxel_generator.VoxelGenerator`` and
``networks.pointpillars8_shared.{PointNet, PointPillarsScatter}`` (nothing is copied), feeds them the
seeded synthetic frames of 3d_object_detection_b200/synth.py and writes

* ``tests/golden/small_<grid>_<case>.npz`` -- inputs + full reference outputs for small cases
  (voxelise, decorate, scatter), and
* ``tests/golden/full_hashes.json``        -- sha256 of inputs and of the reference's voxelisation outputs
  at BASELINE.json's full frame sizes for every grid x {S1,S2,S3,S4}.

/root/reference does not exist on the GPU box, which is why the vectors are committed.
"""
from __future__ import annotations

import hashlib
import importlib
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
GOLD = ROOT / "tests" / "golden"
REF = "/root/reference"

pkg = importlib.import_module("3d_object_detection_b200")
cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")


def sha(a: np.ndarray) -> str:
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.view(np.uint8).reshape(-1).tobytes() if a.size else b"").hexdigest()


def ref_modules():
    sys.path.insert(0, REF)
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    sys.dont_write_bytecode = True
    import torch  # noqa
    from framework.voxel_generator import VoxelGenerator          # the reference's own class
    from networks.pointpillars8_shared import PointNet, PointPillarsScatter
    return VoxelGenerator, PointNet, PointPillarsScatter


def ref_decorate(PointNet, cfg, voxels, num, coors):
    """Run the reference's PointNet.forward with its PFN swapped for a tap, so lines 32-54 execute verbatim
    and we capture the tensor they hand to the PFN."""
    import torch

    class Tap(torch.nn.Module):
        def forward(self, x):
            self.seen = x.detach().clone()
            return x

    net = PointNet(cfg["num_point_features"], cfg["voxel_size"], cfg["detection_offset"])
    net.pfn_layers = Tap()
    with torch.no_grad():
        net(torch.from_numpy(voxels.copy()), torch.from_numpy(num.copy()), torch.from_numpy(coors.copy()))
    return net.pfn_layers.seen.permute(0, 2, 1).contiguous().numpy()      # (V,P,9)


def ref_scatter(PointPillarsScatter, cfg, feats, coors):
    import torch
    sc = PointPillarsScatter(batch_size=1, output_shape=cfg["grid_size"], num_input_features=feats.shape[1])
    with torch.no_grad():
        return sc(torch.from_numpy(feats.copy()), torch.from_numpy(coors.copy())).numpy()


SMALL_CASES = [
    # name, kind, n, n_cells, overrides
    ("s1", "S1", 4000, 400, {}),
    ("s1_overflow", "S1", 4000, 900, {"max_voxels": 300}),
    ("s3_overflow", "S3", 3000, 0, {"max_voxels": 500}),
    ("s2_ring", "S2", 4096, 0, {"max_voxels": 700}),
    ("s4_edges", "S4", 1500, 200, {}),
    ("dense_cap", "S1", 3000, 12, {}),            # few cells, every pillar far over max_num_points
    ("n0", "S1", 0, 10, {}),
    ("n5", "S1", 5, 3, {}),
]


def main():
    VoxelGenerator, PointNet, PointPillarsScatter = ref_modules()
    GOLD.mkdir(parents=True, exist_ok=True)
    full = {}
    for gi, gkey in enumerate(sorted(cfgmod.GRIDS)):
        # ---------- small cases: full outputs ----------
        for ci, (name, kind, n, n_cells, over) in enumerate(SMALL_CASES):
            cfg = cfgmod.builtin_config(gkey)
            cfg.update(over)
            geom = cfgmod.derive_grid(cfg)
            seed = synth.frame_seed(100 + gi, ci)
            pts = synth.make_frame(kind, geom, max(n, 8), seed, max(n_cells, 1))[:n] if n else \
                np.zeros((0, 4), np.float32)
            pts = np.ascontiguousarray(pts, np.float32)
            vg = VoxelGenerator(cfg)                       # mutates cfg (F8)
            assert np.array_equal(vg.grid_size, geom.grid_size) and np.array_equal(vg.offset, geom.offset)
            voxels, coors, num = vg.generate(pts)
            out = dict(points=pts, voxels=voxels, coors=coors, num=num,
                       max_voxels=np.int32(cfg["max_voxels"]), max_points=np.int32(cfg["max_num_points"]),
                       grid_size=vg.grid_size, offset=vg.offset, voxel_size=vg.voxel_size,
                       detection_range=vg.detection_range)
            if len(num):
                out["features"] = ref_decorate(PointNet, cfg, voxels, num, coors)
                if name in ("s1", "s4_edges") and gkey in ("nuscene", "20cm"):
                    rng = np.random.default_rng(seed + 7)
                    feats = rng.standard_normal((len(num), 64)).astype(np.float32)
                    canvas = ref_scatter(PointPillarsScatter, cfg, feats, coors)
                    out["pillar_feats"] = feats
                    out["canvas_sha"] = np.array(sha(canvas))
                    out["canvas_sum"] = np.float64(canvas.astype(np.float64).sum())
                    nz = np.nonzero(canvas.reshape(-1))[0]
                    out["canvas_nz_idx"] = nz.astype(np.int64)
                    out["canvas_nz_val"] = canvas.reshape(-1)[nz]
                    out["canvas_shape"] = np.array(canvas.shape, np.int64)
            np.savez_compressed(GOLD / f"small_{gkey}_{name}.npz", **out)
            print(f"small {gkey:8s} {name:12s} N={len(pts):5d} V={len(num):5d} kept={int(num.sum())}")
        # ---------- full-size frames: hashes ----------
        for ki, kind in enumerate(["S1", "S2", "S3", "S4"]):
            cfg = cfgmod.builtin_config(gkey)
            geom = cfgmod.derive_grid(cfg)
            seed = synth.frame_seed(gi, ki)
            pts = synth.make_frame(kind, geom, cfg["frame_points"], seed, synth.S1_CELLS[gkey])
            vg = VoxelGenerator(cfg)
            voxels, coors, num = vg.generate(pts)
            full[f"{gkey}/{kind}"] = dict(seed=seed, n=int(len(pts)), points_sha=sha(pts), V=int(len(num)),
                                          kept=int(num.sum()), voxels_sha=sha(voxels), coors_sha=sha(coors),
                                          num_sha=sha(num))
            print(f"full  {gkey:8s} {kind} N={len(pts)} V={len(num)} kept={int(num.sum())}")
    (GOLD / "full_hashes.json").write_text(json.dumps(full, indent=1, sort_keys=True) + "\n")


if __name__ == "__main__":
    main()
