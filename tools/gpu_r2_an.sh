#!/bin/bash
# call AN of round 2: default ("auto") PDL = second quantize range of single-frame calls only; whole suite, then A/B on the probe
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2an_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/r2an_tests.txt; tail -3 gpurun_out/r2an_tests.txt
for rep in 1 2; do for pdl in auto 0; do
  if [ $pdl = auto ]; then e=""; else e="PF_PDL=0"; fi
  env $e python tools/frame_probe.py --grid 20cm > gpurun_out/r2an_${pdl}_20cm_$rep.txt 2>&1
  echo "pdl $pdl 20cm rep $rep: $(grep 'generate\|same graph' gpurun_out/r2an_${pdl}_20cm_$rep.txt | tr -s ' ' | tr '\n' ';')"
done; done
for pdl in auto 0; do
  if [ $pdl = auto ]; then e=""; else e="PF_PDL=0"; fi
  env $e python tools/frame_probe.py --grid nuscene > gpurun_out/r2an_${pdl}_nuscene.txt 2>&1
  echo "pdl $pdl nuscene: $(grep 'generate\|same graph' gpurun_out/r2an_${pdl}_nuscene.txt | tr -s ' ' | tr '\n' ';')"
done
python bench.py --no-cpu-baseline --no-pfn --no-anchor-mask --no-postprocess --steps 100 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][-1]); print('bench', round(d['ms_per_step'],4), round(d['e2e']['ms_per_step'],4), {k: round(v,4) for k,v in d['single_frame_api'].items() if isinstance(v,float)})"
