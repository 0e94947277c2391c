#!/bin/bash
# call AH of round 2: host overhead of the stand-alone modules trimmed -- tests that touch them, then the bench's single-frame block
mkdir -p gpurun_out
python -m pytest tests/test_gpu_scatter_pipeline.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2ah_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/r2ah_tests.txt; tail -3 gpurun_out/r2ah_tests.txt
for rep in 1 2; do
python bench.py --no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess --steps 100 > gpurun_out/r2ah_bench_$rep.json 2>/dev/null
python - $rep <<'PY'
import json, sys
d = json.loads([l for l in open(f"gpurun_out/r2ah_bench_{sys.argv[1]}.json") if l.startswith("{")][-1])
print({k: round(v, 4) for k, v in d["single_frame_api"].items() if isinstance(v, float)}, "step %.4f e2e %.4f" % (d["ms_per_step"], d["e2e"]["ms_per_step"]))
PY
done
