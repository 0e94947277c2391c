#!/bin/bash
# round 2, call E: fill variant 6, single-frame phase probe, sharded front end on one GPU, full-set ncu of one eager step
tag=${1:-r2e}
o=gpurun_out
mkdir -p $o
timeout 600 python -m pytest tests/test_gpu_scatter_pipeline.py tests/test_gpu_parity.py -x -q 2>&1 | tail -8 > $o/${tag}_pytest.log
timeout 120 python tools/frame_probe.py > $o/${tag}_frame_probe.txt 2>&1
timeout 120 python tools/frame_probe.py --grid nuscene >> $o/${tag}_frame_probe.txt 2>&1
B="--no-cpu-baseline --no-postprocess --no-anchor-mask --no-single --no-pfn --steps 100"
for v in 1 6; do
  for g in 20cm 10cm nuscene; do
    PF_FILL_VARIANT=$v timeout 300 python bench.py --grid $g $B > $o/${tag}_fill${v}_$g.json 2> $o/${tag}_fill${v}_$g.err
  done
done
timeout 300 python bench.py --sharded --gpus 1 --steps 50 > $o/${tag}_sharded_n1.json 2> $o/${tag}_sharded_n1.err
timeout 120 python tools/prof_step.py --grid 20cm --steps 4 > $o/${tag}_plain_20cm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:^k_ -s 18 -c 8 -f -o $o/${tag}_step_20cm \
    python tools/prof_step.py --grid 20cm --steps 4 > $o/${tag}_ncu_20cm.log 2>&1
cat $o/${tag}_pytest.log; cat $o/${tag}_frame_probe.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2e_*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "e2e ms", round(d["e2e"]["ms_per_step"], 4),
              "fill us", round(1e3 * d.get("stages_ms", {}).get("scatter", 0), 1), "frac", round(d.get("roofline", {}).get("frac", 0), 3))
    except Exception as e:
        print(f, "ERR", e)
PY
