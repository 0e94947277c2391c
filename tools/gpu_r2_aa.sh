#!/bin/bash
# call AA of round 2: host threads of the staging copy inside generate(), same box
mkdir -p gpurun_out
for rep in 1 2; do for thr in 2 4 8; do
  PF_STAGE_THREADS=$thr python tools/frame_probe.py --grid 20cm > gpurun_out/r2aa_t${thr}_$rep.txt 2>&1
  echo "threads $thr rep $rep: $(grep 'generate\|stage threads [248]:' gpurun_out/r2aa_t${thr}_$rep.txt | grep -v side | tr -s ' ' | tr '\n' ';')"
done; done
for thr in 4 8; do
  PF_STAGE_THREADS=$thr python tools/frame_probe.py --grid nuscene > gpurun_out/r2aa_nuscene_t${thr}.txt 2>&1
  echo "nuscene threads $thr: $(grep 'generate\|stage threads [248]:' gpurun_out/r2aa_nuscene_t${thr}.txt | grep -v side | tr -s ' ' | tr '\n' ';')"
done
