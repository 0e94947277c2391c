#!/bin/bash
# round 2, call H: sanity of the sliced column scan + per-slot sharded streams on one GPU, then compute-sanitizer memcheck
tag=${1:-r2h}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests/test_gpu_scatter_pipeline.py tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q 2>&1 | tail -6 > $o/${tag}_pytest.log
for g in 20cm nuscene 10cm; do
  timeout 120 python tools/stage_times.py --grid $g > $o/${tag}_stage_$g.txt 2>&1
done
timeout 300 python bench.py --sharded --gpus 1 --steps 100 > $o/${tag}_sharded_n1.json 2> $o/${tag}_sharded_n1.err
timeout 400 python bench.py --no-cpu-baseline --no-postprocess --no-anchor-mask --no-pfn > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
cat $o/${tag}_pytest.log; cat $o/${tag}_stage_*.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2h_*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "value G", round(d["value"] / 1e9, 3), "e2e ms", round(d["e2e"]["ms_per_step"], 4), d.get("serial"), d.get("single_frame_api"))
    except Exception as e:
        print(f, "ERR", e)
PY
bash tools/gpu_sanitize.sh memcheck
