#!/bin/bash
# round 2, call D: restored chain (ticketed k_assign), STG.256 fill as default, TMA 2 KB variants, single-frame fast path
tag=${1:-r2d}
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
for v in 4 5; do
  for g in 20cm 10cm; do
    PF_FILL_VARIANT=$v timeout 300 python bench.py --grid $g $B > $o/${tag}_fill${v}_$g.json 2> $o/${tag}_fill${v}_$g.err
  done
done
timeout 400 python bench.py --no-cpu-baseline --no-postprocess --no-anchor-mask > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
cat $o/${tag}_pytest.log; cat $o/${tag}_stage_*.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2d_*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "fill us", round(1e3 * d["stages_ms"]["scatter"], 1),
              "frac", round(d["roofline"]["frac"], 3), "vox us", round(1e3 * d["stages_ms"]["voxelize_decorate"], 1), d.get("single_frame_api"))
    except Exception as e:
        print(f, "ERR", e)
PY
