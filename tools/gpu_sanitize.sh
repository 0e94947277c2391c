#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md): bash tools/gpu_sanitize.sh memcheck|racecheck|initcheck|synccheck
tool=${1:-memcheck}
o=gpurun_out
mkdir -p $o
timeout 300 python tools/sanitize_driver.py > $o/sanitize_plain.log 2>&1 || { tail -20 $o/sanitize_plain.log; echo "plain run failed"; exit 1; }
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file $o/sanitize_$tool.log python tools/sanitize_driver.py > $o/sanitize_${tool}_stdout.log 2>&1
echo "exit $?"; tail -5 $o/sanitize_${tool}_stdout.log; tail -15 $o/sanitize_$tool.log
