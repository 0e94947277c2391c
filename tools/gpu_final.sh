#!/bin/bash
# Evidence pass for profiles/: parity tests, bench lines, ncu launch list of the bench command, ncu --set full of one step.
tag=${1:-fin}
o=gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > $o/${tag}_pytest.log
timeout 600 python bench.py > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
for g in nuscene 10cm 16cm; do
  timeout 300 python bench.py --grid $g --no-cpu-baseline > $o/${tag}_bench_$g.json 2> $o/${tag}_bench_$g.err
done
timeout 300 python bench.py --grid 20cm --batch 64 --no-cpu-baseline --no-pfn --no-single > $o/${tag}_bench_20cm_b64.json 2> /dev/null
timeout 300 python bench.py --grid nuscene --batch 64 --no-cpu-baseline --no-pfn --no-single > $o/${tag}_bench_nuscene_b64.json 2> /dev/null
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
# launch list of the bench command itself (plain run first, then the same command under ncu)
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-pfn --no-single > $o/${tag}_ll_plain.json 2> /dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv \
    python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-pfn --no-single > $o/${tag}_ll_ncu.log 2>&1
# full-set capture of one eager step (8 launches at these frame sizes; the 9-launch window also shows the next step's first), after warm-up steps
for g in 20cm nuscene; do
  timeout 120 python tools/prof_step.py --grid $g --steps 4 > $o/${tag}_plain_$g.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:^k_ -s 20 -c 9 -f -o $o/${tag}_step_$g \
      python tools/prof_step.py --grid $g --steps 4 > $o/${tag}_ncu_$g.log 2>&1
done
cat $o/${tag}_pytest.log
