#!/bin/bash
# call AL of round 2: chunked upload in front of the frame graph (shipped) against the upload inside the graph
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_scatter_pipeline.py -m gpu -x -q > gpurun_out/r2al_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/r2al_tests.txt; tail -3 gpurun_out/r2al_tests.txt
PF_FRAME_UPLOAD=graph python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "voxelize_small or generate or pooled" 2>&1 | tail -1
for rep in 1 2; do for mode in chunked graph; do for g in 20cm nuscene 10cm; do
  PF_FRAME_UPLOAD=$mode python tools/frame_probe.py --grid $g > gpurun_out/r2al_${mode}_${g}_$rep.txt 2>&1
  echo "$mode $g rep $rep: $(grep 'generate\|chunked upload' gpurun_out/r2al_${mode}_${g}_$rep.txt | tr -s ' ' | tr '\n' ';')"
done; done; done
