"""GPU probe: how fast can 1.3 GB be written on this B200, by what?  (calibrates the canvas-fill roofline)"""
import importlib, sys, os
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
pf = importlib.import_module("3d_object_detection_b200")

def timeit(fn, n=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3

dev = torch.device("cuda", 0)
B = 8
canvas = torch.empty((B, 64, 800, 800), device=dev)
other = torch.empty_like(canvas)
nbytes = canvas.numel() * 4
t = timeit(lambda: canvas.fill_(0.0));  print(f"torch fill_      {t*1e6:8.1f} us  {nbytes/t/1e9:8.1f} GB/s written")
t = timeit(lambda: canvas.zero_());     print(f"torch zero_      {t*1e6:8.1f} us  {nbytes/t/1e9:8.1f} GB/s written")
t = timeit(lambda: other.copy_(canvas)); print(f"torch copy_      {t*1e6:8.1f} us  {2*nbytes/t/1e9:8.1f} GB/s read+write")
cfg = pf.builtin_config("20cm")
fe = pf.PillarFrontEnd(cfg, max_batch=B, max_points_per_frame=120000, device=0)
pillar = torch.randn((B * 16000, 64), device=dev)
for var in os.environ.get("PF_FILL_VARIANTS", "0").split(","):
    os.environ["PF_FILL_VARIANT"] = var
    t = timeit(lambda: fe.scatter(pillar))
    print(f"pf fill (empty map) variant {var}: {t*1e6:8.1f} us  {nbytes/t/1e9:8.1f} GB/s written")
