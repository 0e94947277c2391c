"""Post-processing probe: time pf_postprocess (PostProcessor) on B frames x 400 x 400 x 9 anchors with CUDA events, eager
and as a CUDA-graph replay, for a sparse and a dense candidate population, and the Inference.infer_gpu drop-in for one
frame (host wall clock incl. the read-back).  Writes one JSON line per measurement.

    python tools/post_probe.py --frames 8 [--out gpurun_out/post_probe.json]
"""
import argparse
import importlib
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
cfgmod = importlib.import_module("3d_object_detection_b200.config")
synth = importlib.import_module("3d_object_detection_b200.synth")
inf = importlib.import_module("3d_object_detection_b200.inference")

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=8)
ap.add_argument("--fmap", type=int, default=400)
ap.add_argument("--reps", type=int, default=50)
ap.add_argument("--out", default="")
a = ap.parse_args()

LIMIT = [-80.0, -80.0, -10.0, 80.0, 80.0, 10.0]
geom = cfgmod.derive_grid(cfgmod.builtin_config("20cm"))
anchors, ranges = synth.anchors_3d(geom, (a.fmap, a.fmap))
A = len(anchors)
lines = []


def emit(**kw):
    lines.append(kw)
    print(json.dumps(kw), flush=True)


def timed(fn, reps):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


post = inf.PostProcessor(anchors, ranges, LIMIT, max_frames=a.frames)
for label, mean_logit in (("sparse", -9.6), ("dense", -6.0), ("all", 3.0)):
    cls, box, dirp, mask = (torch.from_numpy(x).cuda() for x in synth.post_preds(a.frames, A, 7, mean_logit=mean_logit))
    eager = timed(lambda: post(cls, box, dirp, mask), a.reps)
    cand = post.region("cand_count", a.frames).cpu().numpy()[:a.frames * len(ranges)]
    num = post.det_num.cpu().numpy()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        post(cls, box, dirp, mask)
        s.synchronize()
        with torch.cuda.graph(g, stream=s):
            post(cls, box, dirp, mask)
    graph = timed(g.replay, a.reps)
    # bytes the call has to read at least: class logits + anchor mask of every anchor
    emit(what="pf_postprocess", population=label, frames=a.frames, anchors=A, candidates_per_segment=[int(cand.min()), int(cand.max())],
         detections_per_frame=[int(num.min()), int(num.max())], eager_us=round(eager, 1), graph_us=round(graph, 1),
         stream_GBps=round(a.frames * A * 5 / (graph * 1e-6) / 1e9, 1))

cfg = dict(device=0, center_limit=LIMIT, detect_class=["vehicle", "pedestrian", "cyclist"], batch_size=1)


class Assigner:
    pass


Assigner.anchors = anchors
Assigner.class_masks = {n: r for n, r in zip(cfg["detect_class"], ranges)}
I = inf.Inference(cfg, Assigner())
cls, box, dirp, mask = (torch.from_numpy(x).cuda() for x in synth.post_preds(1, A, 8, mean_logit=-9.6))
ex, pr = {"anchors_mask": mask}, {"cls_preds": cls, "box_preds": box, "dir_preds": dirp}
for _ in range(5):
    annos = I.infer_gpu(ex, pr)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(a.reps):
    annos = I.infer_gpu(ex, pr)
emit(what="Inference.infer_gpu drop-in, one frame, host wall clock incl. read-back", anchors=A,
     us_per_call=round(1e6 * (time.perf_counter() - t0) / a.reps, 1), detections=int(len(annos[0]["name"])))
if a.out:
    Path(a.out).write_text("\n".join(json.dumps(x) for x in lines) + "\n")
