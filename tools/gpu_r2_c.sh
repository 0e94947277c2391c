#!/bin/bash
# round 2, call C: k_order (assign + in-tile ranks fused, ticketed) + partial-row column scan; canvas-fill variants A/B
tag=${1:-r2c}
o=gpurun_out
mkdir -p $o
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -15 > $o/${tag}_pytest_small.log
if ! grep -q " passed" $o/${tag}_pytest_small.log || grep -q "failed\|error" $o/${tag}_pytest_small.log; then
  cat $o/${tag}_pytest_small.log; echo "small parity failed -- stopping"; exit 1
fi
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $o/${tag}_pytest.log
for g in 20cm nuscene 10cm; do
  timeout 120 python tools/stage_times.py --grid $g > $o/${tag}_stage_$g.txt 2>&1
done
B="--no-cpu-baseline --no-postprocess --no-anchor-mask --no-single --no-pfn --steps 100"
for v in 0 1 2 3; do
  for g in 20cm nuscene 10cm; do
    PF_FILL_VARIANT=$v timeout 300 python bench.py --grid $g $B > $o/${tag}_fill${v}_$g.json 2> $o/${tag}_fill${v}_$g.err
  done
  PF_FILL_VARIANT=$v timeout 300 python bench.py --grid 20cm --batch 64 $B > $o/${tag}_fill${v}_20cm_b64.json 2> /dev/null
done
cat $o/${tag}_pytest.log; cat $o/${tag}_stage_*.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2c_fill*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "fill us", round(1e3 * d["stages_ms"]["scatter"], 1),
              "frac", round(d["roofline"]["frac"], 3), "vox us", round(1e3 * d["stages_ms"]["voxelize_decorate"], 1))
    except Exception as e:
        print(f, "ERR", e)
PY
