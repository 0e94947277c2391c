#!/bin/bash
# round 2, call B: the fused k_order chain -- small parity first (guards against a hang), then the whole GPU suite,
# stage times and bench lines
tag=${1:-r2b}
o=gpurun_out
mkdir -p $o
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -15 > $o/${tag}_pytest_small.log
if ! grep -q " passed" $o/${tag}_pytest_small.log || grep -q "failed\|error" $o/${tag}_pytest_small.log; then
  cat $o/${tag}_pytest_small.log; echo "small parity failed -- stopping"; exit 1
fi
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $o/${tag}_pytest.log
for g in 20cm nuscene 10cm; do
  timeout 120 python tools/stage_times.py --grid $g > $o/${tag}_stage_$g.txt 2>&1
done
timeout 120 python tools/stage_times.py --grid 20cm --pfn > $o/${tag}_stage_20cm_pfn.txt 2>&1
timeout 400 python bench.py --no-cpu-baseline --no-postprocess --no-anchor-mask > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
timeout 300 python bench.py --grid nuscene --no-cpu-baseline --no-postprocess --no-anchor-mask --no-single > $o/${tag}_bench_nuscene.json 2> $o/${tag}_bench_nuscene.err
timeout 300 python bench.py --grid 10cm --no-cpu-baseline --no-postprocess --no-anchor-mask --no-single --no-pfn > $o/${tag}_bench_10cm.json 2> $o/${tag}_bench_10cm.err
cat $o/${tag}_pytest.log; cat $o/${tag}_stage_*.txt
