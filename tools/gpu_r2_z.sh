#!/bin/bash
# call Z of round 2: whole GPU suite + smoke + default bench line + probe on the tree with the one-call single-frame path
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2z_gpu_tests.txt 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r2z_gpu_tests.txt
tail -4 gpurun_out/r2z_gpu_tests.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2z_smoke.txt 2>&1; tail -1 gpurun_out/r2z_smoke.txt
python bench.py > gpurun_out/r2z_bench_20cm.json 2> gpurun_out/r2z_bench.err
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2z_bench_20cm.json") if l.startswith("{")][-1])
print("value %.3f G pts/s, %.4f ms/step; e2e %.3f G, %.4f ms; fill frac %.3f" % (d["value"] / 1e9, d["ms_per_step"], d["e2e"]["value"] / 1e9, d["e2e"]["ms_per_step"], d["roofline"]["frac"]))
print("single_frame_api", d.get("single_frame_api"))
PY
python tools/frame_probe.py --grid 20cm > gpurun_out/r2z_frame_probe.txt 2>&1; grep "generate\|pf_frame_run" gpurun_out/r2z_frame_probe.txt
python tools/frame_probe.py --grid nuscene > gpurun_out/r2z_frame_probe_nuscene.txt 2>&1; grep "generate\|pf_frame_run" gpurun_out/r2z_frame_probe_nuscene.txt
