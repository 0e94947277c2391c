#!/bin/bash
# BASELINE.json configs[2]-[4] with the shipped kernel set, one bench line each -> profiles/ (run under gpurun):
#   [2] ntusl_10cm: 10 cm grid, 300 k-point frames, batch 32, 1 GPU
#   [3] nuscene: 350 k-point frames, frame-sharded across the GPUs of the box (torchrun, N = visible GPUs when > 1)
#   [4] ntusl_20cm_mly: streaming 64-frame batches, >= 2 s per measurement, all visible GPUs
tag=${1:-r2}
o=gpurun_out
mkdir -p $o
N=$(python -c "import torch; print(torch.cuda.device_count())")
S="--no-cpu-baseline --no-single --no-pfn --no-anchor-mask --no-postprocess"
timeout 600 python bench.py --grid 10cm --batch 32 --min-seconds 2 $S > $o/${tag}_cfg2_10cm_b32_n1.json 2> $o/${tag}_cfg2.err
timeout 600 python bench.py --grid nuscene --batch 8 --min-seconds 2 $S > $o/${tag}_cfg3_nuscene_b8_n1.json 2> $o/${tag}_cfg3.err
timeout 600 python bench.py --grid 20cm --batch 64 --min-seconds 2 $S > $o/${tag}_cfg4_20cm_b64_n1.json 2> $o/${tag}_cfg4.err
if [ "$N" -gt 1 ]; then
  for n in 2 4 8; do
    [ "$n" -le "$N" ] || continue
    R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n))"
    timeout 600 $R bench.py --gpus $n --grid nuscene --batch 8 --min-seconds 2 $S > $o/${tag}_cfg3_nuscene_b8_n$n.json 2>> $o/${tag}_cfg3.err
    timeout 600 $R bench.py --gpus $n --grid 20cm --batch 64 --min-seconds 2 $S > $o/${tag}_cfg4_20cm_b64_n$n.json 2>> $o/${tag}_cfg4.err
  done
fi
ls -la $o/${tag}_cfg*.json
