"""Top stall locations of one kernel from an .ncu-rep (SASS view): python tools/ncu_hot.py rep [N]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(io.StringIO(out)))
# several kernels may follow one another: split on "Kernel Name" rows
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
for b in blocks:
    h = b["rows"][0]; data = b["rows"][1:]
    ci = {x: i for i, x in enumerate(h)}
    tot = sum(int(r[ci["# Samples"]]) for r in data)
    print(f"== {b['name'][:80]}  samples {tot}  sass lines {len(data)}")
    stalls = [x for x in h if x.startswith("stall_") and "Not Issued" not in x]
    agg = {s: sum(int(r[ci[s]]) for r in data) for s in stalls}
    print("   ", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:6]})
    idx = sorted(range(len(data)), key=lambda i: -int(data[i][ci["# Samples"]]))[:n]
    for i in sorted(idx):
        r = data[i]
        top = sorted(((int(r[ci[s]]), s) for s in stalls), reverse=True)[:2]
        print(f"  {i:5d} {int(r[ci['# Samples']]):6d} {100*int(r[ci['# Samples']])/max(tot,1):5.1f}%  {r[ci['Source']].strip()[:70]:70s} {top}")
