#!/bin/bash
o=gpurun_out
rm -f $o/stage.log
for g in 20cm nuscene 10cm; do
  timeout 200 python tools/stage_times.py --grid $g >> $o/stage.log 2>&1
  timeout 200 python tools/stage_times.py --grid $g --pfn >> $o/stage.log 2>&1
done
cat $o/stage.log
