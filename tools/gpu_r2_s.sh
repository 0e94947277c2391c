#!/bin/bash
# call S of round 2: which programmatic-dependent launches pay (PF_PDL_MASK), serial step (one front end)
mkdir -p gpurun_out
SIDE="--no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess --no-single --device-pipelines 1 --steps 200"
for mask in 0x00 0x01 0x03 0x07 0x1f 0x3f 0x40 0x41 0x7f; do
  for cfg in "20cm 8" "nuscene 8" "20cm 1" "10cm 32"; do
    set -- $cfg
    PF_PDL_MASK=$mask python bench.py --grid $1 --batch $2 $SIDE > gpurun_out/r2s_m${mask}_$1_b$2.json 2> gpurun_out/r2s_err.txt
    python - "$mask" "$1" "$2" gpurun_out/r2s_m${mask}_$1_b$2.json <<'PY'
import json, sys
line = [l for l in open(sys.argv[4]) if l.startswith("{")][-1]
d = json.loads(line)
print("mask", sys.argv[1], sys.argv[2], "b", sys.argv[3], "serial ms/step %.4f" % d["ms_per_step"], "fill %.4f" % d["roofline"]["ms_per_launch"], "e2e %.4f" % d["e2e"]["ms_per_step"])
PY
  done
done
