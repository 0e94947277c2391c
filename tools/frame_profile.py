#!/usr/bin/env python
"""Where the host time of the single-frame drop-in call goes: cProfile of VoxelGenerator.generate (numpy in, device out and
numpy out) over many calls on one 20 cm frame.  Host-side diagnosis only; the latencies that count are tools/frame_probe.py's."""
import argparse
import cProfile
import importlib
import io
import pstats
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")
cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")

ap = argparse.ArgumentParser()
ap.add_argument("--grid", default="20cm")
ap.add_argument("--calls", type=int, default=3000)
args = ap.parse_args()
cfg = cfgmod.builtin_config(args.grid)
geom = cfgmod.derive_grid(cfg)
n = cfg["frame_points"]
frame = np.ascontiguousarray(synth.make_frame("S1", geom, n, 5, synth.S1_CELLS[args.grid])[:n], np.float32)
gen = pf.VoxelGenerator(cfg, device=0, max_points_per_frame=n)
for dev_out in (True, False):
    for _ in range(50):
        gen.generate(frame, device_outputs=dev_out)
    t0 = time.perf_counter()
    for _ in range(args.calls):
        out = gen.generate(frame, device_outputs=dev_out)
    dt = (time.perf_counter() - t0) / args.calls * 1e6
    print(f"{args.grid} generate(device_outputs={dev_out}): {dt:.1f} us per call")
    prof = cProfile.Profile()
    prof.enable()
    for _ in range(args.calls):
        out = gen.generate(frame, device_outputs=dev_out)
    prof.disable()
    s = io.StringIO()
    pstats.Stats(prof, stream=s).sort_stats("tottime").print_stats(22)
    print(s.getvalue())
