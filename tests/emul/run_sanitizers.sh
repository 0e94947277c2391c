#!/bin/bash
# TEST INFRASTRUCTURE ONLY.  compute-sanitizer is closed on this GPU pool (gpurun answers "closed ... stays closed"), so the
# memcheck / racecheck role is played on the HOST: the kernel bodies of csrc/pf_voxel_kernels.cuh, compiled with g++ through
# tests/emul/warp_emul.h (one OS thread per CUDA thread), run under AddressSanitizer + UBSan and under ThreadSanitizer and
# are compared with the sequential C oracle.   bash tests/emul/run_sanitizers.sh [cases] > profiles/r2_emul_sanitizers.txt
cd "$(dirname "$0")" && mkdir -p _build
cases=${1:-3}
CS=../../3d_object_detection_b200/csrc
for san in address,undefined thread; do
  name=$(echo $san | tr ',' '_')
  g++ -O1 -g -std=c++17 -ffp-contract=off -pthread -fsanitize=$san -fno-omit-frame-pointer -Wno-unknown-pragmas -Wno-unused-function \
      -I $CS -I . -I ../../include -c voxel_emul.cpp -o _build/ve_$name.o || exit 1
  gcc -O1 -g -ffp-contract=off -fsanitize=$san -fopenmp -c ../../oracle/pillar_oracle.c -o _build/po_$name.o || exit 1
  g++ -O1 -g -std=c++17 -pthread -fsanitize=$san voxel_emul_main.cpp _build/ve_$name.o _build/po_$name.o -o _build/voxel_emul_$name -fopenmp || exit 1
  echo "== -fsanitize=$san : ./voxel_emul_$name $cases"
  TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0" ASAN_OPTIONS="detect_leaks=1" UBSAN_OPTIONS="print_stacktrace=1" \
      ./_build/voxel_emul_$name $cases 2>&1 | grep -E "WARNING|ERROR|SUMMARY|runtime error|case [0-9]|ALL CASES|FAILED|reported" 
  echo "== exit status ${PIPESTATUS[0]}"
done
