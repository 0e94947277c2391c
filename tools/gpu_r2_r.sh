#!/bin/bash
# call R of round 2: programmatic dependent launch of the chain, on / off (PF_PDL), tests first
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2r_gpu_tests.txt 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r2r_gpu_tests.txt
tail -4 gpurun_out/r2r_gpu_tests.txt
SIDE="--no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess"
for pdl in 1 0; do
  for cfg in "20cm 8" "nuscene 8" "20cm 1"; do
    set -- $cfg
    PF_PDL=$pdl python bench.py --grid $1 --batch $2 $SIDE --device-pipelines 1 > gpurun_out/r2r_pdl${pdl}_$1_b$2_dp1.json 2> gpurun_out/r2r_err.txt
    python - "$pdl" "$1" "$2" gpurun_out/r2r_pdl${pdl}_$1_b$2_dp1.json <<'PY'
import json, sys
line = [l for l in open(sys.argv[4]) if l.startswith("{")][-1]
d = json.loads(line)
sf = d.get("single_frame_api", {})
print("pdl", sys.argv[1], sys.argv[2], "b", sys.argv[3], "dp1 ms/step %.4f" % d["ms_per_step"], "fill ms %.4f" % d["roofline"]["ms_per_launch"],
      "e2e ms %.4f" % d["e2e"]["ms_per_step"], {k: v for k, v in sf.items() if isinstance(v, (int, float))})
PY
  done
  PF_PDL=$pdl python bench.py $SIDE > gpurun_out/r2r_pdl${pdl}_20cm_default.json 2>> gpurun_out/r2r_err.txt
  tail -1 gpurun_out/r2r_pdl${pdl}_20cm_default.json | cut -c1-330
  PF_PDL=$pdl python tools/frame_probe.py --grid 20cm > gpurun_out/r2r_pdl${pdl}_frame_probe.txt 2>&1; tail -6 gpurun_out/r2r_pdl${pdl}_frame_probe.txt
done
