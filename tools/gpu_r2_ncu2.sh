#!/bin/bash
# closing full-set ncu captures of one eager step on the final tree (20 cm x 8 and nuScenes x 8); plain run first
o=gpurun_out; mkdir -p $o
for g in 20cm nuscene; do
  timeout 120 python tools/prof_step.py --grid $g --steps 4 > $o/r2n_plain_$g.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:^k_ -s 18 -c 8 -f -o $o/r2n_step_$g \
      python tools/prof_step.py --grid $g --steps 4 > $o/r2n_ncu_$g.log 2>&1
  echo "$g: plain rc, ncu rc $?"; tail -2 $o/r2n_ncu_$g.log
done
ls -la $o/*.ncu-rep
