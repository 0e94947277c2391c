#!/bin/bash
# call AD of round 2: which fill variant on SMALL canvases (one or two frames)
mkdir -p gpurun_out
SIDE="--no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess --no-single --device-pipelines 1 --steps 200"
for cfg in "20cm 1" "20cm 2" "nuscene 1" "nuscene 2" "16cm 1"; do
  set -- $cfg
  for v in 0 1 7 8 6; do
    PF_FILL_VARIANT=$v python bench.py --grid $1 --batch $2 $SIDE > gpurun_out/r2ad_$1_b$2_v$v.json 2>/dev/null
    python - $1 $2 $v <<'PY'
import json, sys
d = json.loads([l for l in open(f"gpurun_out/r2ad_{sys.argv[1]}_b{sys.argv[2]}_v{sys.argv[3]}.json") if l.startswith("{")][-1])
r = d["roofline"]
print(sys.argv[1], "b", sys.argv[2], "variant", sys.argv[3], "fill %.1f us  frac %.3f  step %.1f us" % (1e3 * r["ms_per_launch"], r["frac"], 1e3 * d["ms_per_step"]))
PY
  done
done
