#!/bin/bash
o=gpurun_out
F="--no-cpu-baseline --no-pfn --no-single --no-anchor-mask"
for n in 4 2; do
 for m in spread linear; do
  PF_DEVICE_MAP=$m timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29615 bench.py --gpus $n $F > $o/sc3_${m}_n$n.json 2> $o/sc3_${m}_n$n.err
 done
done
PF_DEVICE_MAP=spread timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29616 bench.py --gpus 4 --grid nuscene $F > $o/sc3_spread_nuscene_n4.json 2> $o/sc3_spread_nuscene_n4.err
tail -c 100 $o/sc3_spread_n4.json
