"""Stall samples per CUDA source line: python tools/ncu_lines.py rep [N]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(io.StringIO(out)))
hdr = None; cur_file = ""; agg = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None: continue
    ci_s = hdr.index("# Samples")
    if r[0] != "":   # a source line header row
        key = (cur_file, int(r[0]), r[1].strip()[:90]); agg.setdefault(key, [0, {}]); last = key
    else:
        try: s = int(r[ci_s])
        except ValueError: continue
        agg[last][0] += s
        for i, h in enumerate(hdr):
            if h.startswith("stall_") and "Not Issued" not in h:
                agg[last][1][h] = agg[last][1].get(h, 0) + int(r[i])
tot = sum(v[0] for v in agg.values())
print("total samples", tot)
for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    top = sorted(v[1].items(), key=lambda kv: -kv[1])[:3]
    print(f"{key[0]}:{key[1]:4d} {v[0]:6d} {100*v[0]/max(tot,1):5.1f}%  {key[2]:90s} {[(a[6:], b) for a, b in top]}")
