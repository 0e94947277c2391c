#!/bin/bash
# call X of round 2: does the OpenMP wait policy of the staging copy explain the slow replays after it?
mkdir -p gpurun_out
nproc; cat /sys/fs/cgroup/cpu.max 2>/dev/null; python -c "import os; print('affinity', len(os.sched_getaffinity(0)))"
for pol in default passive spin0; do
  case $pol in
    default) env_pref="";;
    passive) env_pref="OMP_WAIT_POLICY=passive";;
    spin0) env_pref="GOMP_SPINCOUNT=0";;
  esac
  echo "== $pol"
  env $env_pref python tools/frame_probe.py --grid 20cm > gpurun_out/r2x_frame_probe_$pol.txt 2>&1
  grep -v "pf_stage_copy 1.92 MB, [128]\|np.copyto\|clones\|D2H of V" gpurun_out/r2x_frame_probe_$pol.txt
done
