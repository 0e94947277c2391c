#!/bin/bash
# Round-2 closing evidence pass for profiles/ (one GPU, final tree): GPU suite, smoke, bench lines of every workload, the
# reference arm, stage times, single-frame probes, the ncu launch list of the bench command and one full-set capture of an
# eager step (each ncu run only after the same command exited 0 without it).
tag=r2fin2
o=gpurun_out
mkdir -p $o
timeout 1800 python -m pytest tests -m gpu -q > $o/${tag}_pytest.log 2>&1; echo "suite rc=$?" >> $o/${tag}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $o/${tag}_smoke.txt 2>&1; echo "smoke rc=$?" >> $o/${tag}_smoke.txt
timeout 900 python bench.py > $o/${tag}_bench_20cm.json 2> $o/${tag}_bench_20cm.err
S="--no-cpu-baseline --no-anchor-mask --no-postprocess"
for g in nuscene 10cm 16cm; do
  timeout 400 python bench.py --grid $g $S > $o/${tag}_bench_$g.json 2> $o/${tag}_bench_$g.err
done
timeout 400 python bench.py --grid 20cm --batch 64 $S --no-pfn --no-single > $o/${tag}_bench_20cm_b64.json 2> /dev/null
timeout 400 python bench.py --device-pipelines 1 $S --no-pfn --no-single > $o/${tag}_bench_20cm_dp1.json 2> /dev/null
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
for g in 20cm nuscene 10cm; do
  timeout 120 python tools/stage_times.py --grid $g > $o/${tag}_stage_$g.txt 2>&1
  timeout 120 python tools/stage_times.py --grid $g --pfn >> $o/${tag}_stage_$g.txt 2>&1
done
timeout 120 python tools/frame_probe.py > $o/${tag}_frame_probe.txt 2>&1
timeout 120 python tools/frame_probe.py --grid nuscene >> $o/${tag}_frame_probe.txt 2>&1
L="--steps 30 --warmup 5 --no-cpu-baseline --no-pfn --no-single --no-anchor-mask --no-postprocess"
timeout 300 python bench.py $L > $o/${tag}_ll_plain.json 2> /dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/${tag}_launches.csv \
    python bench.py $L > $o/${tag}_ll_ncu.log 2>&1
tail -3 $o/${tag}_pytest.log; tail -2 $o/${tag}_smoke.txt
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2fin2_*.json")):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 4), "G", round(d["value"] / 1e9, 3), "serial", round(d.get("serial", {}).get("ms_per_step", 0), 4),
              "e2e ms", round(d["e2e"].get("ms_per_step", 0), 4), "fill frac", round(d.get("roofline", {}).get("frac", 0), 3), "step frac", round(d.get("e2e_hbm_frac", 0), 3))
    except Exception as e:
        print(f, "ERR", e)
PY
