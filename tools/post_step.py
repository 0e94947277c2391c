"""A few pf_postprocess calls on one population, for ncu (launch list / --set full).
    python tools/post_step.py --population dense --reps 3"""
import argparse
import importlib
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")
inf = importlib.import_module("3d_object_detection_b200.inference")
ap = argparse.ArgumentParser()
ap.add_argument("--population", default="dense", choices=["sparse", "dense", "all"])
ap.add_argument("--frames", type=int, default=8)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
geom = cfgmod.derive_grid(cfgmod.builtin_config("20cm"))
anchors, ranges = synth.anchors_3d(geom, (400, 400))
post = inf.PostProcessor(anchors, ranges, [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0], max_frames=a.frames)
mean = {"sparse": -9.6, "dense": -6.0, "all": 3.0}[a.population]
ins = [torch.from_numpy(x).cuda() for x in synth.post_preds(a.frames, len(anchors), 7, mean_logit=mean)]
for _ in range(a.reps):
    post(*ins)
torch.cuda.synchronize()
print(a.population, post.det_num.cpu().tolist())
