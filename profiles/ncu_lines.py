"""Per-source-line dynamic warp-instruction counts from `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`."""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
fil = None; hdr = None; agg = defaultdict(lambda: [0, 0, ""])
for r in rows:
    if not r: continue
    if r[0] == "File Path": fil = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or len(r) < 8: continue
    ln = r[0]
    if r[2] == "-" and ln:       # a CUDA source line row: aggregated metrics for that line
        try:
            agg[(fil, int(ln))][0] += int(r[hdr["Instructions Executed"]] or 0)
            agg[(fil, int(ln))][1] += int(r[hdr["# Samples"]] or 0)
            agg[(fil, int(ln))][2] = r[1].strip()[:90]
        except ValueError:
            pass
tot = sum(v[0] for v in agg.values())
it = int(sys.argv[2]) if len(sys.argv) > 2 else 1
print("total warp instructions", tot, "per iteration", round(tot / it, 1))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[3]) if len(sys.argv) > 3 else 45]:
    print(f"{k[0]:>28}:{k[1]:<4} {v[0] / it:7.2f}  samples {v[1]:6d}  {v[2]}")
