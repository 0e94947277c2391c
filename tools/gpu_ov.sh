#!/bin/bash
o=gpurun_out
for g in 20cm nuscene 10cm; do
  timeout 200 python tools/overlap_probe.py --grid $g >> $o/ov.log 2>&1
  timeout 200 python tools/overlap_probe.py --grid $g --same-priority >> $o/ov.log 2>&1
  timeout 200 python tools/overlap_probe.py --grid $g --depth 3 >> $o/ov.log 2>&1
done
cat $o/ov.log
