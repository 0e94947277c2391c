#!/bin/bash
# call Y of round 2: staging copy with non-temporal stores (default) against plain memcpy (PF_STAGE_NT=0)
mkdir -p gpurun_out
for nt in 1 0 1 0; do
  echo "== PF_STAGE_NT=$nt"
  PF_STAGE_NT=$nt python tools/frame_probe.py --grid 20cm > gpurun_out/r2y_frame_probe_nt${nt}.txt 2>&1
  grep -v "pf_stage_copy 1.92 MB, [128]\|np.copyto\|clones" gpurun_out/r2y_frame_probe_nt${nt}.txt
done
for nt in 1 0; do
  echo "== nuscene PF_STAGE_NT=$nt"
  PF_STAGE_NT=$nt python tools/frame_probe.py --grid nuscene > gpurun_out/r2y_frame_probe_nuscene_nt${nt}.txt 2>&1
  grep "pf_stage_copy\|generate\|same graph" gpurun_out/r2y_frame_probe_nuscene_nt${nt}.txt
done
