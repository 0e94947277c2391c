#!/bin/bash
# call AE of round 2: small-canvas fill rule -- GPU suite, single-frame probe, default bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2ae_gpu_tests.txt 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r2ae_gpu_tests.txt
tail -3 gpurun_out/r2ae_gpu_tests.txt
python tools/frame_probe.py --grid 20cm > gpurun_out/r2ae_frame_probe.txt 2>&1; grep "generate" gpurun_out/r2ae_frame_probe.txt
python tools/frame_probe.py --grid nuscene > gpurun_out/r2ae_frame_probe_nuscene.txt 2>&1; grep "generate" gpurun_out/r2ae_frame_probe_nuscene.txt
python bench.py > gpurun_out/r2ae_bench_20cm.json 2> gpurun_out/r2ae_bench.err
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2ae_bench_20cm.json") if l.startswith("{")][-1])
print("value %.3f G pts/s, %.4f ms/step; e2e %.3f G, %.4f ms; fill frac %.3f" % (d["value"] / 1e9, d["ms_per_step"], d["e2e"]["value"] / 1e9, d["e2e"]["ms_per_step"], d["roofline"]["frac"]))
print({k: round(v, 4) for k, v in d["single_frame_api"].items() if isinstance(v, float)})
PY
