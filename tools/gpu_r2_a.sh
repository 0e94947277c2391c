#!/bin/bash
# round 2, call A: the round-1 kernel set on this round's box -- GPU suite incl. the new every-frame full-size tests,
# baseline bench lines, stage times, and the copy-only host-link probe at N = 1
tag=${1:-r2a}
o=gpurun_out
mkdir -p $o
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $o/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $o/${tag}_pytest.log
timeout 400 python bench.py > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
timeout 300 python bench.py --grid nuscene --no-cpu-baseline --no-postprocess --no-anchor-mask > $o/${tag}_bench_nuscene.json 2> $o/${tag}_bench_nuscene.err
timeout 120 python tools/stage_times.py > $o/${tag}_stage_times.txt 2>&1
timeout 120 python tools/h2d_ceiling.py --ns 1 --out $o/${tag}_h2d_n1.json > $o/${tag}_h2d.log 2>&1
cat $o/${tag}_pytest.log
