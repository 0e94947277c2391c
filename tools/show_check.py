"""Print the essentials of a tools/gpu_check.sh run: python tools/show_check.py <tag>"""
import csv, json, sys
tag = sys.argv[1]
print(open(f"gpurun_out/{tag}_pytest.log").read().strip().splitlines()[-1])
for g in ["20cm", "nuscene", "10cm"]:
    try:
        d = json.loads(open(f"gpurun_out/{tag}_bench_{g}.json").read())
        print(f"{g:8s} ms/step {d['ms_per_step']:.4f}  fps {d['frames_per_s']:.0f}  stages {d['stages_ms']}  "
              f"fill frac {d['roofline']['frac']:.3f}  e2e_hbm_frac {d['e2e_hbm_frac']:.3f}  e2e ms {d['e2e']['ms_per_step']:.3f}")
    except Exception as e:
        print(g, "bench failed:", e, open(f"gpurun_out/{tag}_bench_{g}.err").read()[-800:])
for g in ["20cm", "nuscene"]:
    try:
        rows = [r for r in csv.reader(open(f"gpurun_out/{tag}_ll_{g}.csv")) if len(r) > 10]
    except Exception as e:
        print(g, "no launch list", e); continue
    h = rows[0]; d = {}
    for r in rows[1:]:
        k = (int(r[0]), r[h.index("Kernel Name")].split("(")[0][:28])
        d.setdefault(k, {})[r[h.index("Metric Name")]] = float(r[h.index("Metric Value")])
    print(f"-- {g} launch list (ncu, cold, serialised)")
    tot = 0
    items = sorted(d.items())
    starts = [i for i, (k, v) in enumerate(items) if "k_quantize" in k[1] and (i == 0 or "k_quantize" not in items[i - 1][0][1])]
    lo = starts[-2] if len(starts) >= 2 and len(items) - starts[-1] < starts[-1] - starts[-2] else starts[-1]
    hi = lo + (starts[-1] - starts[-2] if len(starts) >= 2 else len(items) - lo)
    for k, v in items[lo:hi]:
        tot += v["gpu__time_duration.sum"]
        print(f"  {k[1]:28s} {v['gpu__time_duration.sum']/1e3:8.1f} us  rd {v['dram__bytes_read.sum']/1e6:7.1f} MB  "
              f"wr {v['dram__bytes_write.sum']/1e6:7.1f} MB  inst {v['smsp__inst_executed.sum']/1e6:6.2f} M")
    print(f"  total {tot/1e3:.1f} us")
