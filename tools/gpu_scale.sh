#!/bin/bash
# 8-GPU box: topology, then the e2e arm at N = 8 and 4 with and without the NUMA-local host binding
o=gpurun_out
(nvidia-smi topo -m; lscpu | grep -i -E "numa|^CPU\(s\)|model name|socket"; nproc; cat /sys/fs/cgroup/cpuset.cpus.effective 2>/dev/null) > $o/sc_topo.txt 2>&1
F="--no-cpu-baseline --no-pfn --no-single --no-anchor-mask"
for n in 8 4; do
  for aff in "" "--no-affinity"; do
    timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $n $F $aff > $o/sc_n${n}${aff}.json 2> $o/sc_n${n}${aff}.err
  done
done
cat $o/sc_topo.txt
