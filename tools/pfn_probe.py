"""GPU probe: whole step with the learned PFN in the loop -- fused (pf_voxelize_pfn) vs unfused (CUDA decoration +
torch Conv1d/BatchNorm1d/ReLU/max) vs the PFN-less benchmark step (static pillar features)."""
import importlib, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")
bench = importlib.import_module("bench")
grid = sys.argv[1] if len(sys.argv) > 1 else "20cm"
cfg = pf.builtin_config(grid); cfg["_name"] = grid
geom = pf.derive_grid(cfg)
B = 8
frames, pts_cat, offs = bench.make_batch(pf, cfg, geom, B, 0)
dev = torch.device("cuda", 0)
pts = torch.from_numpy(pts_cat).to(dev); offs_d = torch.from_numpy(offs).to(dev)
fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0)
net = pf.PointNet(4, cfg["voxel_size"], geom.offset).to(dev).eval()
pillar = torch.randn((B * cfg["max_voxels"], 64), device=dev)

def timeit(fn, label, n=50):
    with torch.no_grad():
        for _ in range(5): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"{grid} {label:58s} {ms:8.3f} ms/step  {B / ms * 1e3:8.0f} frames/s")

timeit(lambda: fe.step(pts, offs_d, pillar_feats=pillar), "no PFN (static pillar features; the bench step)")
timeit(lambda: fe.step(pts, offs_d, pfn=net), "fused PFN (pf_voxelize_pfn + fill)")
def unfused():
    fe.voxelize_decorate(pts, offs_d, keep_cell_map=False)
    fe.scatter(net.features_to_pillars(fe.features))
timeit(unfused, "unfused PFN (decorate -> torch conv/bn/relu/max -> fill)")
timeit(lambda: fe.voxelize_pfn(pts, offs_d, net.folded()), "  voxelize + decorate + PFN only (fused)")
timeit(lambda: fe.voxelize_decorate(pts, offs_d, keep_cell_map=False), "  voxelize + decorate only")
