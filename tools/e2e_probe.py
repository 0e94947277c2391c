"""GPU probe: end-to-end (pinned host points -> canvas on device, coords/counts back) step time against the
number of pipelines in flight, repeated, to see how stable the copy/compute overlap is."""
import importlib, sys, time
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
offs_d = torch.from_numpy(offs).to(dev)
pillar = torch.randn((B * cfg["max_voxels"], 64), device=dev)
K = 60

def bw_probe():
    h = torch.from_numpy(pts_cat).pin_memory(); d = torch.empty_like(h, device=dev)
    for _ in range(3): d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): d.copy_(h, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"H2D {h.numel()*4/1e6:.1f} MB pinned: {ms*1e3:.0f} us = {h.numel()*4/ms/1e6:.1f} GB/s")

bw_probe()
for depth in (1, 2, 3, 4):
    pipes = []
    for d in range(depth):
        fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=cfg["frame_points"], device=0)
        pipes.append(dict(fe=fe, stream=torch.cuda.Stream(dev), h_pts=torch.from_numpy(pts_cat).pin_memory(),
                          d_pts=torch.empty((pts_cat.shape[0], 4), device=dev),
                          h_cnt=torch.empty((B,), dtype=torch.int32).pin_memory(),
                          h_coors=torch.empty((B * cfg["max_voxels"], 4), dtype=torch.int32).pin_memory()))
    def step(i):
        p = pipes[i % depth]
        with torch.cuda.stream(p["stream"]):
            p["d_pts"].copy_(p["h_pts"], non_blocking=True)
            p["fe"].step(p["d_pts"], offs_d, pillar_feats=pillar)
            p["h_cnt"].copy_(p["fe"].voxel_num, non_blocking=True)
            p["h_coors"].copy_(p["fe"].coors, non_blocking=True)
    res = []
    for rep in range(4):
        for i in range(6): step(i)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cur = torch.cuda.current_stream(dev)
        e0.record(cur)
        for p in pipes: p["stream"].wait_event(e0)
        for i in range(K): step(i)
        t_host = time.perf_counter() - t0
        for p in pipes: cur.wait_stream(p["stream"])
        e1.record(cur); torch.cuda.synchronize()
        res.append((round(e0.elapsed_time(e1) / K, 4), round(1e3 * t_host / K, 4)))
    print(f"depth {depth}: (gpu ms/step, host enqueue ms/step) x4 = {res}")
    for p in pipes: p["fe"].close()
