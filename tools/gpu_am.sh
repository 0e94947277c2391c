#!/bin/bash
# anchor-mask check: parity tests, bench side measurement, launch list of the four kernels
tag=${1:-am}
o=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $o/${tag}_pytest.log
for g in 20cm 10cm nuscene; do
  timeout 120 python tools/am_probe.py --grid $g > $o/${tag}_probe_$g.log 2>&1 && \
  timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,smsp__inst_executed.sum --clock-control none \
      -k regex:^k_am --csv --log-file $o/${tag}_ll_$g.csv python tools/am_probe.py --grid $g --reps 2 > /dev/null 2>&1
done
cat $o/${tag}_pytest.log
