"""GPU probe: how much of the voxelise chain is launch/dependency overhead?  Graph of pf_voxelize_decorate only,
replayed on real frames and on empty frames (every kernel exits at once)."""
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
pts = torch.from_numpy(pts_cat).to(dev)
fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0)

def timed(offs_np, label, eager=False):
    offs_d = torch.from_numpy(offs_np).to(dev)
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        for _ in range(3): fe.voxelize_decorate(pts, offs_d, keep_cell_map=False)
    torch.cuda.synchronize()
    if not eager:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fe.voxelize_decorate(pts, offs_d, keep_cell_map=False)
        run = g.replay
    else:
        run = lambda: fe.voxelize_decorate(pts, offs_d, keep_cell_map=False)
    for _ in range(5): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200): run()
    e1.record(); torch.cuda.synchronize()
    print(f"{grid} {label}: {e0.elapsed_time(e1) / 200 * 1e3:.1f} us per voxelize_decorate")

timed(offs, "real frames, graph")
timed(offs, "real frames, eager", eager=True)
timed(np.zeros_like(offs), "empty frames, graph")
small = (offs // 64).astype(np.int32)
timed(small, "1/64-size frames, graph")
