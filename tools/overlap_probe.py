"""Does the voxelise chain of batch k+1 hide under the canvas fill of batch k?  Two front ends (double buffering),
chains on a high-priority stream, fills on a low-priority one, events between them.  Eager launches.

    python tools/overlap_probe.py --grid 20cm [--same-priority]
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
ap.add_argument("--steps", type=int, default=200)
ap.add_argument("--depth", type=int, default=2)
ap.add_argument("--same-priority", action="store_true")
a = ap.parse_args()

cfg = pf.builtin_config(a.grid)
cfg["_name"] = a.grid
geom = pf.derive_grid(cfg)
B = a.batch or (32 if "10cm" in a.grid else 8)
frames, pts_cat, offs = bench.make_batch(pf, cfg, geom, B, 0)
dev = torch.device("cuda", 0)
fes = [pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0) for _ in range(a.depth)]
pts = torch.from_numpy(pts_cat).to(dev)
offs_d = torch.from_numpy(offs).to(dev)
pillar = torch.from_numpy(np.random.default_rng(1).standard_normal((B * cfg["max_voxels"], 64)).astype(np.float32)).to(dev)
lo_p, hi_p = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
s_chain = torch.cuda.Stream(dev, priority=0 if a.same_priority else -1)
s_fill = torch.cuda.Stream(dev, priority=0)
torch.cuda.synchronize()


def run(n):
    done_fill = [None] * a.depth
    for k in range(n):
        fe = fes[k % a.depth]
        with torch.cuda.stream(s_chain):
            if done_fill[k % a.depth] is not None:
                s_chain.wait_event(done_fill[k % a.depth])      # workspace / outputs of this front end are free again
            fe.voxelize_decorate(pts, offs_d)
            e = torch.cuda.Event()
            e.record(s_chain)
        with torch.cuda.stream(s_fill):
            s_fill.wait_event(e)
            fe.scatter(pillar)
            f = torch.cuda.Event()
            f.record(s_fill)
            done_fill[k % a.depth] = f


run(20)
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record(s_chain)
s_fill.wait_event(t0)
run(a.steps)
s_fill.wait_stream(s_chain)
t1.record(s_fill)
torch.cuda.synchronize()
ms = t0.elapsed_time(t1) / a.steps
# serial reference on one stream
fe = fes[0]
for _ in range(10):
    fe.step(pts, offs_d, pillar_feats=pillar)
torch.cuda.synchronize()
u0, u1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
u0.record()
for _ in range(a.steps):
    fe.step(pts, offs_d, pillar_feats=pillar)
u1.record()
torch.cuda.synchronize()
print(f"{a.grid} batch {B}: overlapped {ms:.4f} ms/step (depth {a.depth}, chain priority {'same' if a.same_priority else 'high'}), "
      f"serial eager {u0.elapsed_time(u1) / a.steps:.4f} ms/step, voxels {fes[0].voxel_num.cpu().tolist()[:4]}")
