#!/bin/bash
# round 2, call I (one GPU): sanity of the current tree before the 8-GPU call
tag=${1:-r2i}
o=gpurun_out
mkdir -p $o
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > $o/${tag}_pytest.log
timeout 300 python bench.py --sharded --gpus 1 --steps 100 > $o/${tag}_sharded_n1.json 2> $o/${tag}_sharded_n1.err
timeout 300 python bench.py --sharded --pinned-frames --gpus 1 --steps 100 > $o/${tag}_sharded_pinned_n1.json 2> $o/${tag}_sharded_pinned_n1.err
timeout 120 python tools/stage_times.py --grid 20cm > $o/${tag}_stage_20cm.txt 2>&1
cat $o/${tag}_pytest.log $o/${tag}_stage_20cm.txt; tail -3 $o/${tag}_sharded*.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2i_*.json")):
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "value G", round(d["value"] / 1e9, 3), d["detail"].get("max_worker_graph_ms_per_step"))
    except Exception as e:
        print(f, "ERR", e)
PY
