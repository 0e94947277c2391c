#!/bin/bash
# call P of round 2: the allocation-free scatter test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2p_noalloc_alone.txt 2>&1; echo "alone rc=$?" | tee -a gpurun_out/r2p_noalloc_alone.txt
tail -3 gpurun_out/r2p_noalloc_alone.txt
python -m pytest tests -m gpu -x -q > gpurun_out/r2p_gpu_tests.txt 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r2p_gpu_tests.txt
tail -5 gpurun_out/r2p_gpu_tests.txt
