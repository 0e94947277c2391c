#!/bin/bash
# closing validation of the round's last tree: GPU suite, smoke, the default bench line, the reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2close_tests.txt 2>&1; echo "suite rc=$?" >> gpurun_out/r2close_tests.txt; tail -3 gpurun_out/r2close_tests.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2close_smoke.txt 2>&1; tail -1 gpurun_out/r2close_smoke.txt
python bench.py > gpurun_out/r2close_bench.json 2> gpurun_out/r2close_bench.err; python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r2close_bench.json") if l.startswith("{")][-1])
print("value %.3f G pts/s, %.4f ms/step; e2e %.3f G, %.4f ms; fill frac %.3f; launches %s" % (d["value"] / 1e9, d["ms_per_step"], d["e2e"]["value"] / 1e9, d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["gpu_launches"]))
print(d["roofline"]["kernel"]); print({k: round(v, 4) for k, v in d["single_frame_api"].items() if isinstance(v, float)})
PY
python bench.py --impl reference --steps 10 --warmup 3 2>/dev/null | tail -1 | cut -c1-250
