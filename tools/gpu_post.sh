#!/bin/bash
# the post-processing row on the GPU, most valuable first (the call may be cut by the remaining GPU budget):
# its parity tests, smoke(), the probe, one bench line, the whole GPU suite, then ncu launch lists of the step
o=gpurun_out
mkdir -p $o
timeout 160 python -m pytest tests/test_gpu_post.py -q > $o/post_pytest.log 2>&1
echo "post tests exit $?" >> $o/post_pytest.log
tail -30 $o/post_pytest.log
timeout 50 python -c "import __graft_entry__ as g; g.smoke()" > $o/smoke.log 2>&1
echo "smoke exit $?" >> $o/smoke.log
tail -3 $o/smoke.log
timeout 60 python tools/post_probe.py --out $o/post_probe.json > $o/post_probe.log 2>&1
tail -6 $o/post_probe.log
timeout 120 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-single --no-pfn > $o/bench_post.json 2> $o/bench_post.err
tail -c 600 $o/bench_post.json
timeout 240 python -m pytest tests -m gpu -x -q > $o/all_pytest.log 2>&1
echo "all tests exit $?" >> $o/all_pytest.log
tail -5 $o/all_pytest.log
for p in sparse dense; do
  timeout 60 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_post|k_nms' --csv \
    --log-file $o/post_launches_$p.csv python tools/post_step.py --population $p --reps 3 > $o/post_ncu_$p.log 2>&1
done
