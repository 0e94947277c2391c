#!/bin/bash
# call AJ of round 2: generate_batch from a list of numpy frames through one staged upload -- tests, then old against new timing
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_anchor_mask.py tests/test_gpu_scatter_pipeline.py -m gpu -x -q > gpurun_out/r2aj_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/r2aj_tests.txt; tail -3 gpurun_out/r2aj_tests.txt
python - <<'PY'
import importlib, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
pf = importlib.import_module("3d_object_detection_b200")
cfgmod = importlib.import_module("3d_object_detection_b200.config"); synth = importlib.import_module("3d_object_detection_b200.synth")
cfg = cfgmod.builtin_config("20cm"); geom = cfgmod.derive_grid(cfg); n = cfg["frame_points"]
frames = [np.ascontiguousarray(synth.make_frame("S1", geom, n, 5 + i, synth.S1_CELLS["20cm"])[:n], np.float32) for i in range(8)]
vg = pf.VoxelGenerator(cfg, device=0, max_points_per_frame=n)
def t(fn, reps=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return 1e3 * (time.perf_counter() - t0) / reps
print("generate_batch(8 numpy frames, device outputs): %.3f ms (staged upload)" % t(lambda: vg.generate_batch(frames)))
devs = lambda: vg._run([vg._to_device(f)[0] for f in frames], compact=True, coors_stride=4)[3].cpu()
print("the previous path (pageable .to() per frame):     %.3f ms" % t(devs))
print("8 x generate(device_outputs=True):                %.3f ms" % t(lambda: [vg.generate(f, device_outputs=True) for f in frames]))
PY
