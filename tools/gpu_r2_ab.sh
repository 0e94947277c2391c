#!/bin/bash
# call AB of round 2: the bench's single_frame_api block, staging with plain memcpy + 4 threads (the previous state) against
# streaming stores + 8 threads (shipped), same box, twice
mkdir -p gpurun_out
SIDE="--no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess --steps 100"
for rep in 1 2; do for arm in old new; do
  if [ $arm = old ]; then e="PF_STAGE_NT=0 PF_STAGE_THREADS=4"; else e="PF_STAGE_NT=1"; fi
  env $e python bench.py $SIDE > gpurun_out/r2ab_${arm}_$rep.json 2>/dev/null
  python - $arm $rep <<'PY'
import json, sys
d = json.loads([l for l in open(f"gpurun_out/r2ab_{sys.argv[1]}_{sys.argv[2]}.json") if l.startswith("{")][-1])
s = d["single_frame_api"]
print(sys.argv[1], sys.argv[2], {k: round(v, 4) for k, v in s.items() if isinstance(v, float)}, "step %.4f e2e %.4f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]))
PY
done; done
