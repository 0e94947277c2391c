"""Where the time of one single-frame VoxelGenerator.generate goes (host wall clock, us): staging copy, graph replay + wait,
read-back / clones.  python tools/frame_probe.py [--grid 20cm]"""
import argparse
import importlib
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")
_lib = importlib.import_module("3d_object_detection_b200._lib")
bench = importlib.import_module("bench")

ap = argparse.ArgumentParser()
ap.add_argument("--grid", default="20cm")
a = ap.parse_args()
cfg, geom, _ = bench.load_workload(pf, a.grid, 1)
frame = bench.make_batch(pf, cfg, geom, 1, 0)[0][0]
vg = pf.VoxelGenerator(dict(cfg), device=0, max_points_per_frame=cfg["frame_points"])
for _ in range(5):
    vg.generate(frame)
    vg.generate(frame, device_outputs=True)
path = next(iter(vg._paths.values()))
lib = _lib.load()
n = len(frame)


def t(fn, reps=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return 1e6 * (time.perf_counter() - t0) / reps


rows = min(path.cap, max(4096, (n + 4095) // 4096 * 4096))
g, g_exec = path.graph_for(rows)
lib.pf_stage_copy(path.np_pts.ctypes.data, frame.ctypes.data, frame.nbytes, 4)
path.np_offs[0], path.np_offs[1] = 0, n
print(f"{a.grid}: frame of {n} points, staged rows {rows}, host threads {vg._threads}")
for thr in (1, 2, 4, 8):
    print(f"  pf_stage_copy {frame.nbytes / 1e6:.2f} MB, {thr} threads: {t(lambda: lib.pf_stage_copy(path.np_pts.ctypes.data, frame.ctypes.data, frame.nbytes, thr)):7.1f} us")
print(f"  np.copyto into pinned:               {t(lambda: np.copyto(path.np_pts[:n], frame)):7.1f} us")


def replay_wait():
    with torch.cuda.stream(path.stream):
        g.replay()
        path.event.record(path.stream)
    path.event.synchronize()


print(f"  graph replay (H2D + 7 kernels + D2H of V) + event wait: {t(replay_wait):7.1f} us")
import ctypes as C  # noqa: E402
c_v = C.c_int32(0)
raw = torch.cuda.current_stream(path.dev).cuda_stream
side = path.stream.cuda_stream
g0, g0_exec = path.graph_for(0)
print(f"  pf_frame_run, chunked upload in front of the graph, {vg._threads} threads: "
      f"{t(lambda: lib.pf_frame_run(path.dev.index, g0_exec, raw, path.ev_done, None, path.p_stage, frame.ctypes.data, frame.nbytes, vg._threads, path.p_dpts, path.p_offs, n, path.p_vnum, C.byref(c_v))):7.1f} us")
for name, st in (("current (default) stream", raw), ("side stream", side)):
    for thr in (0, vg._threads):
        nb = frame.nbytes if thr else 0
        print(f"  pf_frame_run on the {name}, stage threads {thr}: "
              f"{t(lambda: lib.pf_frame_run(path.dev.index, g_exec, st, path.ev_done, None, path.p_stage, frame.ctypes.data, nb, max(thr, 1), None, path.p_offs, n, path.p_vnum, C.byref(c_v))):7.1f} us")


def replay_wait_default():
    g.replay()
    path.event.record(torch.cuda.current_stream(path.dev))
    path.event.synchronize()


print(f"  torch replay + event wait on the current (default) stream: {t(replay_wait_default):7.1f} us")
print(f"  torch replay + event wait on the side stream again:        {t(replay_wait):7.1f} us")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(path.stream):
    e0.record(path.stream)
    for _ in range(100):
        g.replay()
    e1.record(path.stream)
e1.synchronize()
print(f"  the same graph on the device (CUDA events, back to back): {10 * e0.elapsed_time(e1):7.1f} us")
v = int(path.np_vnum[0])
print(f"  3 device clones of V = {v} rows: {t(lambda: (path.voxels[:v].clone(), path.coors[:v].clone(), path.num[:v].clone())):7.1f} us")
blk = vg._out_block()


def d2h():
    with torch.cuda.stream(path.stream):
        blk.tv[:v].copy_(path.voxels[:v], non_blocking=True)
        blk.tc[:v].copy_(path.coors[:v], non_blocking=True)
        blk.tn[:v].copy_(path.num[:v], non_blocking=True)
        path.stream.synchronize()


print(f"  D2H of V rows into a pinned block + stream sync: {t(d2h):7.1f} us")
print(f"  generate(numpy) -> numpy:  {t(lambda: vg.generate(frame)):7.1f} us")
print(f"  generate(numpy) -> device: {t(lambda: vg.generate(frame, device_outputs=True)):7.1f} us")
net = pf.PointNet(4, cfg["voxel_size"], geom.offset).cuda().eval()
with torch.no_grad():
    print(f"  generate_and_scatter:      {t(lambda: vg.generate_and_scatter(frame, net)):7.1f} us")
