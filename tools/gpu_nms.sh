#!/bin/bash
tag=${1:-nms}
o=gpurun_out
timeout 600 python -m pytest tests/test_gpu_nms.py -m gpu -x -q 2>&1 | tail -15 > $o/${tag}_pytest.log
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 >> $o/${tag}_pytest.log
timeout 120 python tools/nms_probe.py > $o/${tag}_probe.log 2>&1
timeout 120 python tools/nms_probe.py --segments 1 >> $o/${tag}_probe.log 2>&1
timeout 120 python tools/nms_probe.py --n 4096 --segments 8 >> $o/${tag}_probe.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:^k_nms --csv \
   --log-file $o/${tag}_ll.csv python tools/nms_probe.py --reps 2 > /dev/null 2>&1
cat $o/${tag}_pytest.log $o/${tag}_probe.log
