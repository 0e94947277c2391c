#!/bin/bash
# round 2, call K (one GPU): fused-PFN gather with 12-float staged rows; quantise split and pipeline-depth experiments
tag=${1:-r2k}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests/test_gpu_scatter_pipeline.py tests/test_gpu_full_size.py -x -q 2>&1 | tail -6 > $o/${tag}_pytest.log
for g in 20cm nuscene 10cm; do
  timeout 120 python tools/stage_times.py --grid $g --pfn > $o/${tag}_stage_${g}_pfn.txt 2>&1
done
for d in 0 8 32; do
  PF_QUANTIZE_SPLIT=$d timeout 120 python tools/stage_times.py --grid 20cm > $o/${tag}_stage_20cm_qsplit$d.txt 2>&1
  PF_QUANTIZE_SPLIT=$d timeout 120 python tools/stage_times.py --grid nuscene > $o/${tag}_stage_nuscene_qsplit$d.txt 2>&1
done
B="--no-cpu-baseline --no-postprocess --no-anchor-mask --no-single --steps 200"
timeout 300 python bench.py $B --device-pipelines 3 > $o/${tag}_dp3_20cm.json 2> /dev/null
timeout 300 python bench.py $B --grid nuscene --device-pipelines 3 > $o/${tag}_dp3_nuscene.json 2> /dev/null
timeout 300 python bench.py $B > $o/${tag}_dp2_20cm.json 2> /dev/null
cat $o/${tag}_pytest.log; grep -h "batch\|quantize\|gather\|sum" $o/${tag}_stage_*.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2k_*.json")):
    d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
    print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "G", round(d["value"] / 1e9, 3), d.get("with_pfn"))
PY
