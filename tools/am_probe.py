"""Anchor-mask probe: voxelise one batch, then a few pf_anchor_mask calls (for ncu launch lists / --set full).

    python tools/am_probe.py --grid 20cm
    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
        -k regex:^k_am --csv --log-file out.csv python tools/am_probe.py --grid 20cm
"""
import argparse
import importlib
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")
synth = importlib.import_module("3d_object_detection_b200.synth")
amod = importlib.import_module("3d_object_detection_b200.anchor_mask")
bench = importlib.import_module("bench")

ap = argparse.ArgumentParser()
ap.add_argument("--grid", default="20cm")
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--reps", type=int, default=4)
a = ap.parse_args()
cfg = pf.builtin_config(a.grid)
cfg["_name"] = a.grid
geom = pf.derive_grid(cfg)
B = a.batch or (32 if "10cm" in a.grid else 8)
frames, pts_cat, offs = bench.make_batch(pf, cfg, geom, B, 0)
dev = torch.device("cuda", 0)
fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0, want_voxels=False)
fe.voxelize_decorate(torch.from_numpy(pts_cat).to(dev), torch.from_numpy(offs).to(dev), keep_cell_map=False)
fmap = (int(geom.grid_size[0]) // 2, int(geom.grid_size[1]) // 2)
am = amod.AnchorMask(synth.anchors_bv(geom, fmap), voxel_size=np.asarray(cfg["voxel_size"], np.float32),
                     offset=geom.offset, grid_size=geom.grid_size)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.reps + 1)]
ev[0].record()
for i in range(a.reps):
    m = fe.anchor_mask(am)
    ev[i + 1].record()
torch.cuda.synchronize()
print("anchor mask ms:", [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(a.reps)], "anchors", am.n_anchors,
      "set", float(m.float().mean()))
