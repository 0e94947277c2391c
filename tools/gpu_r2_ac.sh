#!/bin/bash
# call AC of round 2 (8 GPUs): ShardedFrontEnd from pageable frames with the streaming-store staging copy, against plain memcpy
o=gpurun_out; mkdir -p $o
show() { python - "$1" <<'PY'
import json, sys
d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print(sys.argv[1], "%.2f G pts/s, %.3f ms/step" % (d["value"] / 1e9, d["ms_per_step"]))
PY
}
for nt in 1 0; do for thr in 1 2; do
  PF_STAGE_NT=$nt timeout 300 python bench.py --sharded --gpus 8 --steps 100 --stage-threads $thr > $o/r2ac_sharded_n8_nt${nt}_t$thr.json 2> $o/r2ac_err.txt
  show $o/r2ac_sharded_n8_nt${nt}_t$thr.json
done; done
timeout 300 python bench.py --sharded --pinned-frames --gpus 8 --steps 100 > $o/r2ac_sharded_pinned_n8.json 2>> $o/r2ac_err.txt; show $o/r2ac_sharded_pinned_n8.json
PF_STAGE_NT=1 timeout 300 python bench.py --sharded --gpus 1 --steps 100 > $o/r2ac_sharded_n1_nt1.json 2>> $o/r2ac_err.txt; show $o/r2ac_sharded_n1_nt1.json
PF_STAGE_NT=0 timeout 300 python bench.py --sharded --gpus 1 --steps 100 > $o/r2ac_sharded_n1_nt0.json 2>> $o/r2ac_err.txt; show $o/r2ac_sharded_n1_nt0.json
