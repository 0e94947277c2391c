"""Per-kernel times INSIDE a running step (pf_debug_stage_timing): 5 eager steps back to back, the CUDA events of the
last one read out, repeated; unlike ncu nothing is flushed or serialised, so each kernel sees the cache state the
previous kernel and the previous step's fill really leave behind.

    python tools/stage_times.py --grid 20cm [--pfn]
"""
import argparse
import ctypes as C
import importlib
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")
_lib = importlib.import_module("3d_object_detection_b200._lib")
bench = importlib.import_module("bench")

ap = argparse.ArgumentParser()
ap.add_argument("--grid", default="20cm")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--samples", type=int, default=30)
ap.add_argument("--pfn", action="store_true", help="fused PFN in the gather kernel instead of the decorated rows")
a = ap.parse_args()
cfg = pf.builtin_config(a.grid)
cfg["_name"] = a.grid
geom = pf.derive_grid(cfg)
B = a.batch or (32 if "10cm" in a.grid else 8)
frames, pts_cat, offs = bench.make_batch(pf, cfg, geom, B, 0)
dev = torch.device("cuda", 0)
fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0)
pts = torch.from_numpy(pts_cat).to(dev)
offs_d = torch.from_numpy(offs).to(dev)
pillar = torch.from_numpy(np.random.default_rng(1).standard_normal((B * cfg["max_voxels"], 64)).astype(np.float32)).to(dev)
net = pf.PointNet(4, cfg["voxel_size"], geom.offset).to(dev).eval() if a.pfn else None
lib = fe.handle.lib
_lib.check(lib.pf_debug_stage_timing(fe.handle.ptr, 1))
names = ["quantize(x2)", "assign", "tile_rank", "column_scan", "slot", "gather", "gap", "canvas_fill"]
NS = len(names)
buf = (C.c_float * NS)()
acc = []
with torch.no_grad():
    for s in range(a.samples + 3):
        for _ in range(5):
            if net is not None:
                fe.step(pts, offs_d, pfn=net)
            else:
                fe.step(pts, offs_d, pillar_feats=pillar)
        n = lib.pf_debug_stage_times(fe.handle.ptr, buf, NS)
        assert n == NS
        if s >= 3:
            acc.append([max(buf[i], 0.0) * 1e3 for i in range(NS)])
acc = np.asarray(acc)
print(f"{a.grid} batch {B}{' fused PFN' if a.pfn else ''}: per-kernel us inside the step (mean / min of {a.samples} samples)")
for i, nm in enumerate(names):
    print(f"  {nm:14s} {acc[:, i].mean():8.1f} {acc[:, i].min():8.1f}")
print(f"  {'sum':14s} {acc.sum(1).mean():8.1f} {acc.sum(1).min():8.1f}")
