#!/bin/bash
# parity tests + the three bench lines + in-step per-kernel times
tag=${1:-q}
o=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $o/${tag}_pytest.log
F="--steps 200 --warmup 10 --no-cpu-baseline --no-single --no-anchor-mask --no-pfn"
for g in 20cm nuscene 10cm; do
  timeout 300 python bench.py --grid $g $F > $o/${tag}_bench_$g.json 2> $o/${tag}_bench_$g.err
  timeout 200 python tools/stage_times.py --grid $g >> $o/${tag}_stage.log 2>&1
done
cat $o/${tag}_pytest.log $o/${tag}_stage.log
