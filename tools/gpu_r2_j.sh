#!/bin/bash
# round 2, call J (one GPU): new single-frame tests, smoke(), the restored column scan at 10 cm
tag=${1:-r2j}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_scatter_pipeline.py -x -q 2>&1 | tail -8 > $o/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" > $o/${tag}_smoke.log 2>&1
timeout 120 python tools/stage_times.py --grid 10cm > $o/${tag}_stage_10cm.txt 2>&1
cat $o/${tag}_pytest.log; tail -3 $o/${tag}_smoke.log; cat $o/${tag}_stage_10cm.txt
