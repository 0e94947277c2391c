#!/bin/bash
# quick GPU check: parity tests, bench lines (20 cm / nuScenes / 10 cm) and per-kernel launch lists
tag=${1:-chk}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/${tag}_pytest.log
for g in 20cm nuscene 10cm; do
  timeout 300 python bench.py --grid $g --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_$g.json 2> gpurun_out/${tag}_bench_$g.err
done
for g in 20cm nuscene; do
  timeout 120 python tools/prof_step.py --grid $g --steps 4 > gpurun_out/${tag}_plain_$g.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none \
      -k regex:^k_ -s 20 -c 18 --csv --log-file gpurun_out/${tag}_ll_$g.csv python tools/prof_step.py --grid $g --steps 4 > /dev/null 2>&1
done
cat gpurun_out/${tag}_pytest.log
