#!/bin/bash
# BASELINE.json configs[3] (nuScenes frame-sharded over 2/4/8 GPUs) and configs[4] (streaming 64-frame batches, 8 GPUs)
o=gpurun_out
F="--no-cpu-baseline --no-pfn --no-single --no-anchor-mask"
for n in 2 4 8; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $n --grid nuscene $F > $o/sc2_nuscene_n$n.json 2> $o/sc2_nuscene_n$n.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29614 bench.py --gpus 8 --grid 20cm --batch 64 --steps 100 $F > $o/sc2_20cm_b64_n8.json 2> $o/sc2_20cm_b64_n8.err
tail -c 200 $o/sc2_20cm_b64_n8.json
