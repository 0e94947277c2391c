#!/bin/bash
# round 2, call G (8 GPUs): copy-only host-link ceiling at N = 1, 2, 4, 8; the sharded package runner on every GPU;
# torchrun scaling lines; BASELINE configs[3]/[4] at N = 8
tag=${1:-r2g}
o=gpurun_out
mkdir -p $o
nvidia-smi topo -m > $o/${tag}_topo.txt 2>&1
timeout 300 python tools/h2d_ceiling.py --ns 1,2,4,8 --out $o/${tag}_h2d_ceiling.json > $o/${tag}_h2d.log 2>&1
timeout 200 python tools/h2d_ceiling.py --ns 8 --d2h-mb 2.05 --out $o/${tag}_h2d_ceiling_d2h.json >> $o/${tag}_h2d.log 2>&1
timeout 200 python tools/h2d_ceiling.py --ns 4 --placement linear --out $o/${tag}_h2d_ceiling_linear4.json >> $o/${tag}_h2d.log 2>&1
timeout 300 python -m pytest tests/test_gpu_scatter_pipeline.py -q -k sharded 2>&1 | tail -3 > $o/${tag}_pytest_sharded.log
S="--no-cpu-baseline --no-single --no-pfn --no-anchor-mask --no-postprocess"
for n in 8 4 2; do
  R="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n))"
  timeout 300 $R bench.py --gpus $n --steps 100 $S > $o/${tag}_scale_20cm_n$n.json 2> $o/${tag}_scale_20cm_n$n.err
done
timeout 300 python bench.py --sharded --gpus 8 --steps 100 > $o/${tag}_sharded_n8.json 2> $o/${tag}_sharded_n8.err
timeout 300 python bench.py --sharded --pinned-frames --gpus 8 --steps 100 > $o/${tag}_sharded_pinned_n8.json 2> $o/${tag}_sharded_pinned_n8.err
timeout 300 python bench.py --sharded --pinned-frames --gpus 1 --steps 100 > $o/${tag}_sharded_pinned_n1.json 2> $o/${tag}_sharded_pinned_n1.err
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29790"
timeout 400 $R bench.py --gpus 8 --grid nuscene --batch 8 --min-seconds 2 $S > $o/${tag}_cfg3_nuscene_b8_n8.json 2> $o/${tag}_cfg3.err
timeout 400 $R bench.py --gpus 8 --grid 20cm --batch 64 --min-seconds 2 $S > $o/${tag}_cfg4_20cm_b64_n8.json 2> $o/${tag}_cfg4.err
cat $o/${tag}_h2d.log; cat $o/${tag}_pytest_sharded.log
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2g_*.json")):
    if "h2d" in f: continue
    try:
        d = json.load(open(f))
        print(f.split("/")[-1], "N", d["n_gpus"], "ms/step", round(d["ms_per_step"], 4), "value G", round(d["value"] / 1e9, 3), "e2e ms", round(d["e2e"]["ms_per_step"], 4),
              "e2e G", round(d["e2e"]["value"] / 1e9, 3), d["e2e"].get("roofline"))
    except Exception as e:
        print(f, "ERR", e)
PY
