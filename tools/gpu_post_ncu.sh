#!/bin/bash
# ncu of the post-processing step: launch lists (sparse, dense) and one --set full capture of the dense step
o=gpurun_out
mkdir -p $o
for p in sparse dense; do
  timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_post|k_nms' --csv \
    --log-file $o/post_launches_$p.csv python tools/post_step.py --population $p --reps 3 > $o/post_ncu_$p.log 2>&1
done
timeout 150 ncu --set full --clock-control none --import-source on -k regex:'k_post|k_nms' -s 7 -c 7 \
  -o $o/post_dense_full python tools/post_step.py --population dense --reps 2 > $o/post_ncu_full.log 2>&1
ls -la $o | tail -8
