#!/bin/bash
# round 2, call N (one GPU): full GPU suite + bench lines with the back-to-back fill figure
tag=${1:-r2n}
o=gpurun_out
mkdir -p $o
timeout 1800 python -m pytest tests -m gpu -q 2>&1 | tail -6 > $o/${tag}_pytest.log
timeout 900 python bench.py > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
S="--no-cpu-baseline --no-anchor-mask --no-postprocess --no-single --no-pfn"
for g in nuscene 10cm; do
  timeout 400 python bench.py --grid $g $S > $o/${tag}_bench_$g.json 2> $o/${tag}_bench_$g.err
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" > $o/${tag}_smoke.log 2>&1
cat $o/${tag}_pytest.log; tail -2 $o/${tag}_smoke.log; tail -3 $o/${tag}_bench_20cm.err
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2n_*.json")):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "G", round(d["value"] / 1e9, 3), "dp", d.get("device_pipelines"), "fill", round(d["roofline"]["frac"], 3), d["roofline"].get("back_to_back"))
    except Exception as e:
        print(f, "ERR", e)
PY
