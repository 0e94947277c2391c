#!/bin/bash
# round 2, call L (one GPU): native runner behind ShardedFrontEnd, over-long frame test, automatic pipeline depth
tag=${1:-r2l}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests/test_gpu_scatter_pipeline.py tests/test_gpu_parity.py -x -q 2>&1 | tail -8 > $o/${tag}_pytest.log
timeout 300 python bench.py --sharded --gpus 1 --steps 200 > $o/${tag}_sharded_n1.json 2> $o/${tag}_sharded_n1.err
timeout 300 python bench.py --sharded --pinned-frames --gpus 1 --steps 200 > $o/${tag}_sharded_pinned_n1.json 2> $o/${tag}_sharded_pinned_n1.err
B="--no-cpu-baseline --no-postprocess --no-anchor-mask --no-single --no-pfn --steps 200"
timeout 300 python bench.py $B --grid nuscene > $o/${tag}_bench_nuscene.json 2> /dev/null
timeout 300 python bench.py $B > $o/${tag}_bench_20cm.json 2> /dev/null
cat $o/${tag}_pytest.log; tail -n 3 $o/${tag}_sharded_n1.err $o/${tag}_sharded_pinned_n1.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2l_*.json")):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "G", round(d["value"] / 1e9, 3), d.get("device_pipelines"), d["detail"].get("max_worker_graph_ms_per_step"))
    except Exception as e:
        print(f, "ERR", e)
PY
