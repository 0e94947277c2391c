#!/bin/bash
# round 2, call M (8 GPUs): ShardedFrontEnd on native runners at N = 8 / 4, pageable and pinned frames; parity on every GPU
tag=${1:-r2m}
o=gpurun_out
mkdir -p $o
timeout 300 python -m pytest tests/test_gpu_scatter_pipeline.py -q -k sharded 2>&1 | tail -3 > $o/${tag}_pytest_sharded.log
for n in 8 4; do
  timeout 300 python bench.py --sharded --gpus $n --steps 200 > $o/${tag}_sharded_n$n.json 2> $o/${tag}_sharded_n$n.err
  timeout 300 python bench.py --sharded --pinned-frames --gpus $n --steps 200 > $o/${tag}_sharded_pinned_n$n.json 2> $o/${tag}_sharded_pinned_n$n.err
done
timeout 300 python bench.py --sharded --pinned-frames --gpus 8 --grid nuscene --steps 200 > $o/${tag}_sharded_pinned_nuscene_n8.json 2> /dev/null
cat $o/${tag}_pytest_sharded.log; tail -n 2 $o/${tag}_sharded_n8.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2m_*.json")):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f.split("/")[-1], "N", d["n_gpus"], "ms/step", round(d["ms_per_step"], 4), "G", round(d["value"] / 1e9, 3), d["detail"].get("max_worker_graph_ms_per_step"))
    except Exception as e:
        print(f, "ERR", e)
PY
