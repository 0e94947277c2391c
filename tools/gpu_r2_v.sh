#!/bin/bash
# call V of round 2: single-frame path through pf_frame_run / pf_frame_copy_out -- tests that touch it, then the probes
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_scatter_pipeline.py -m gpu -x -q > gpurun_out/r2v_tests.txt 2>&1; echo "rc=$?" >> gpurun_out/r2v_tests.txt; tail -4 gpurun_out/r2v_tests.txt
python tools/frame_probe.py --grid 20cm > gpurun_out/r2v_frame_probe.txt 2>&1; cat gpurun_out/r2v_frame_probe.txt
python tools/frame_probe.py --grid nuscene >> gpurun_out/r2v_frame_probe.txt 2>&1; tail -14 gpurun_out/r2v_frame_probe.txt
python tools/frame_profile.py --calls 2000 > gpurun_out/r2v_frame_profile.txt 2>&1; grep "us per call" gpurun_out/r2v_frame_profile.txt; head -20 gpurun_out/r2v_frame_profile.txt
