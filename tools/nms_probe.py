"""NMS probe: time pf_nms on S segments of n boxes (CUDA events) and the single-call drop-in nms() (host wall clock).

    python tools/nms_probe.py --n 1000 --segments 24
"""
import argparse
import importlib
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
synth = importlib.import_module("3d_object_detection_b200.synth")
nms = importlib.import_module("3d_object_detection_b200.nms")

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1000)
ap.add_argument("--segments", type=int, default=24)      # 3 classes x 8 frames
ap.add_argument("--thresh", type=float, default=0.1)
ap.add_argument("--reps", type=int, default=50)
a = ap.parse_args()
dets = [synth.nms_dets(a.n, 100 + i) for i in range(a.segments)]
cat = np.concatenate(dets)
offs = torch.from_numpy(np.arange(a.segments + 1, dtype=np.int32) * a.n).cuda()
boxes = torch.from_numpy(cat[:, :4].copy()).cuda()
scores = torch.from_numpy(cat[:, 4].copy()).cuda()
for _ in range(3):
    keep, num = nms.nms_device(boxes, scores, offs, a.n, a.thresh, 300)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.reps):
    keep, num = nms.nms_device(boxes, scores, offs, a.n, a.thresh, 300, keep, num)
e1.record()
torch.cuda.synchronize()
print(f"pf_nms: {a.segments} segments x {a.n} boxes: {1e3 * e0.elapsed_time(e1) / a.reps:.1f} us per batch, kept {num.cpu().tolist()[:6]}")
d0 = dets[0]
for _ in range(3):
    nms.nms(d0[:, :4], d0[:, 4], 1000, 300, a.thresh)
t0 = time.perf_counter()
for _ in range(a.reps):
    nms.nms(d0[:, :4], d0[:, 4], 1000, 300, a.thresh)
print(f"nms() drop-in, numpy in / numpy out, one segment of {a.n}: {1e6 * (time.perf_counter() - t0) / a.reps:.0f} us per call (host wall clock)")
