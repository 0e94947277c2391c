"""Profiling driver: a few EAGER steps of the front end (no CUDA graph, no CPU baseline) so that ncu sees one
launch per kernel per step.  One step = 7 voxelise/decorate kernels + 1 canvas fill (PillarFrontEnd.step).

    python tools/prof_step.py --grid 20cm --steps 4
    ncu --set full --clock-control none --import-source on -k regex:pf::k_ -s <2 + 8*skip> -c 8 -o out \
        python tools/prof_step.py --grid 20cm --steps 4          # (2 = the k_fill_i32 launches of pf_create)
"""
import argparse
import importlib
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")
bench = importlib.import_module("bench")

ap = argparse.ArgumentParser()
ap.add_argument("--grid", default="20cm")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--no-voxels", action="store_true")
a = ap.parse_args()

cfg = pf.builtin_config(a.grid)
cfg["_name"] = a.grid
geom = pf.derive_grid(cfg)
B = a.batch or (32 if "10cm" in a.grid else 8)
frames, pts_cat, offs = bench.make_batch(pf, cfg, geom, B, 0)
dev = torch.device("cuda", 0)
fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0,
                       want_voxels=not a.no_voxels)
pts = torch.from_numpy(pts_cat).to(dev)
offs_d = torch.from_numpy(offs).to(dev)
pillar = torch.from_numpy(np.random.default_rng(1).standard_normal((B * cfg["max_voxels"], 64)).astype(np.float32)).to(dev)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
ev[0].record()
for i in range(a.steps):
    fe.step(pts, offs_d, pillar_feats=pillar)
    ev[i + 1].record()
torch.cuda.synchronize()
print("step ms:", [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(a.steps)],
      "voxels:", fe.voxel_num.cpu().tolist())
