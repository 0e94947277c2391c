#!/bin/bash
# call AF of round 2: fill variants with TWO front ends in flight (the bench default) -- does a leaner fill CTA leave more
# room for the other front end's chain?
mkdir -p gpurun_out
SIDE="--no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess --no-single --steps 300"
for rep in 1 2; do for v in 1 7 8 0; do
  for dp in 2 3; do
    PF_FILL_VARIANT=$v python bench.py $SIDE --device-pipelines $dp > gpurun_out/r2af_v${v}_dp${dp}_$rep.json 2>/dev/null
    python - $v $dp $rep <<'PY'
import json, sys
d = json.loads([l for l in open(f"gpurun_out/r2af_v{sys.argv[1]}_dp{sys.argv[2]}_{sys.argv[3]}.json") if l.startswith("{")][-1])
print("variant", sys.argv[1], "pipelines", sys.argv[2], "rep", sys.argv[3], "step %.4f ms  serial %.4f  fill %.1f us  e2e %.4f" % (d["ms_per_step"], d["serial"]["ms_per_step"], 1e3 * d["roofline"]["ms_per_launch"], d["e2e"]["ms_per_step"]))
PY
  done
done; done
