#!/bin/bash
# ncu --set full of the widened rows: anchor mask (N2) and NMS (N4) kernels
o=gpurun_out
timeout 120 python tools/am_probe.py --grid 20cm --reps 2 > $o/rows_am_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:^k_am -s 4 -c 4 -f -o $o/rows_am \
    python tools/am_probe.py --grid 20cm --reps 2 > $o/rows_am_ncu.log 2>&1
timeout 120 python tools/nms_probe.py --reps 2 > $o/rows_nms_plain.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on -k regex:^k_nms -s 9 -c 3 -f -o $o/rows_nms \
    python tools/nms_probe.py --reps 2 > $o/rows_nms_ncu.log 2>&1
tail -2 $o/rows_am_ncu.log $o/rows_nms_ncu.log
