#!/bin/bash
# round 2, call F: full GPU suite on the current tree, single-frame probe after the hand-out rework, auto fill selection,
# bench lines of the headline + batch-64 config, the sharded runner on one GPU
tag=${1:-r2f}
o=gpurun_out
mkdir -p $o
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $o/${tag}_pytest.log
timeout 120 python tools/frame_probe.py > $o/${tag}_frame_probe.txt 2>&1
B="--no-cpu-baseline --no-postprocess --no-anchor-mask --no-single --no-pfn --steps 100"
for v in 1 6; do
  PF_FILL_VARIANT=$v timeout 300 python bench.py --grid 20cm --batch 64 $B > $o/${tag}_fill${v}_20cm_b64.json 2> /dev/null
done
for g in 20cm nuscene 10cm; do
  timeout 300 python bench.py --grid $g $B --device-pipelines 2 > $o/${tag}_dp2_$g.json 2> $o/${tag}_dp2_$g.err
  timeout 300 python bench.py --grid $g $B --device-pipelines 2 --chain-priority > $o/${tag}_dp2prio_$g.json 2> $o/${tag}_dp2prio_$g.err
done
timeout 300 python bench.py --sharded --gpus 1 --steps 50 > $o/${tag}_sharded_n1.json 2> $o/${tag}_sharded_n1.err
timeout 600 python bench.py > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
cat $o/${tag}_pytest.log; cat $o/${tag}_frame_probe.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2f_*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "value G", round(d["value"] / 1e9, 3), "e2e ms", round(d["e2e"]["ms_per_step"], 4),
              "fill us", round(1e3 * d.get("stages_ms", {}).get("scatter", 0), 1), "frac", round(d.get("roofline", {}).get("frac", 0), 3), d.get("single_frame_api"))
    except Exception as e:
        print(f, "ERR", e)
PY
